// wide_bvh.cpp -- host-side collapse of a Compact2 BVH into the private 8-wide compressed form
// described in ../wide_bvh.h.  Pure host C++ (no CUDA); one-off work inside b2rt_set_bvh.
#include "../wide_bvh.h"

#include <math.h>
#include <string.h>

#include <algorithm>
#include <deque>

namespace b2rt {

namespace {

struct Box { float lo[3], hi[3]; };

inline float area(const Box& b)
{
    const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

// One child of a wide node while it is being assembled.
struct Child {
    Box box;
    int inner;          // >= 0: binary node index to collapse further; -1: leaf part
    int rowStart;       // leaf part: first Woop row of its first triangle (caller's array)
    int count;          // leaf part: 1..3 triangles
};

struct Binary {
    const float* nodes; size_t numNodes;
    const float* woop; size_t numRows;

    Box childBox(int node, int c) const
    {
        const float* p = nodes + (size_t)node * 16;
        Box b;
        if (c == 0) { b.lo[0] = p[0]; b.hi[0] = p[1]; b.lo[1] = p[2]; b.hi[1] = p[3]; b.lo[2] = p[8]; b.hi[2] = p[9]; }
        else        { b.lo[0] = p[4]; b.hi[0] = p[5]; b.lo[1] = p[6]; b.hi[1] = p[7]; b.lo[2] = p[10]; b.hi[2] = p[11]; }
        return b;
    }
    int childRef(int node, int c) const
    {
        int32_t v;
        memcpy(&v, nodes + (size_t)node * 16 + 12 + c, 4);
        return v;
    }
    // number of triangles of the leaf starting at `row` (rows until the 0x80000000 terminator), -1 if malformed
    int leafCount(int row) const
    {
        int n = 0;
        for (size_t r = (size_t)row;; r += 3, n++) {
            if (r >= numRows) return -1;
            uint32_t bits;
            memcpy(&bits, woop + r * 4, 4);
            if (bits == 0x80000000u) return n;
            if (r + 2 >= numRows) return -1;
        }
    }
};

// Children of binary node `node`, leaves cut into parts of at most three triangles (same box).
bool appendChildren(const Binary& bin, int node, std::vector<Child>& out)
{
    for (int c = 0; c < 2; c++) {
        const int ref = bin.childRef(node, c);
        Child ch;
        ch.box = bin.childBox(node, c);
        if (ref >= 0) {
            if (ref % 4 != 0 || (size_t)(ref / 4) >= bin.numNodes) return false;
            ch.inner = ref / 4; ch.rowStart = 0; ch.count = 0;
            out.push_back(ch);
        } else {
            const int row = ~ref;
            const int n = bin.leafCount(row);
            if (n < 0) return false;
            for (int first = 0; first < n; first += 3) {
                ch.inner = -1; ch.rowStart = row + first * 3; ch.count = std::min(3, n - first);
                out.push_back(ch);
            }
        }
    }
    return true;
}

int slotsOfBinaryChild(const Binary& bin, int node, int c)
{
    const int ref = bin.childRef(node, c);
    if (ref >= 0) return 1;
    const int n = bin.leafCount(~ref);
    return n <= 0 ? (n < 0 ? 99 : 0) : (n + 2) / 3;
}

}  // namespace

bool buildWide8(const float* nodes2, size_t numNodes2, const float* woop, const int32_t* triIndex, size_t numRows,
                Wide8& out, const char** error)
{
    static const char* kBad = "malformed Compact2 arrays (child index or leaf run out of range)";
    Binary bin{ nodes2, numNodes2, woop, numRows };
    out = Wide8();
    if (numNodes2 == 0) { if (error) *error = "empty node array"; return false; }

    struct Pending { int binaryNode; int depth; };
    std::deque<Pending> queue;
    queue.push_back(Pending{ 0, 1 });
    out.nodes.resize(20, 0u);                               // wide node 0 = root
    size_t next = 0;                                        // wide node being written
    double childSum = 0.0;

    std::vector<Child> ch;
    while (!queue.empty()) {
        const Pending cur = queue.front(); queue.pop_front();
        // ---- collapse: open the inner child with the largest box while the result fits eight slots ----
        ch.clear();
        if (!appendChildren(bin, cur.binaryNode, ch)) { if (error) *error = kBad; return false; }
        while (true) {
            int best = -1; float bestArea = -1.0f;
            for (size_t i = 0; i < ch.size(); i++) {
                if (ch[i].inner < 0) continue;
                const int need = slotsOfBinaryChild(bin, ch[i].inner, 0) + slotsOfBinaryChild(bin, ch[i].inner, 1);
                if ((int)ch.size() - 1 + need > 8) continue;
                const float a = area(ch[i].box);
                if (a > bestArea) { bestArea = a; best = (int)i; }
            }
            if (best < 0) break;
            const int open = ch[best].inner;
            ch.erase(ch.begin() + best);
            if (!appendChildren(bin, open, ch)) { if (error) *error = kBad; return false; }
        }
        if (ch.size() > 8) { if (error) *error = "internal: wide node overflow"; return false; }
        childSum += (double)ch.size();
        if (cur.depth > out.maxDepth) out.maxDepth = cur.depth;

        // ---- node frame: origin and per-axis power-of-two grid step ----
        Box nb; for (int k = 0; k < 3; k++) { nb.lo[k] = 3.0e38f; nb.hi[k] = -3.0e38f; }
        for (const Child& c : ch) for (int k = 0; k < 3; k++) { nb.lo[k] = std::min(nb.lo[k], c.box.lo[k]); nb.hi[k] = std::max(nb.hi[k], c.box.hi[k]); }
        if (ch.empty()) for (int k = 0; k < 3; k++) nb.lo[k] = nb.hi[k] = 0.0f;
        int e[3];
        for (int k = 0; k < 3; k++) {
            const double ext = (double)nb.hi[k] - (double)nb.lo[k];
            if (!(ext > 0.0)) { e[k] = -141; continue; }
            int ek = (int)ceil(log2(ext / 255.0));
            while (ceil(ext / ldexp(1.0, ek)) > 255.0) ek++;
            e[k] = std::max(-141, std::min(112, ek));
        }

        // ---- slot assignment: slot bit 4/2/1 set <=> child lies on the +x/+y/+z side, so that
        //      (slot XOR ray octant) orders children front to back (greedy auction on the centroid offsets) ----
        int slotOf[8]; bool slotUsed[8] = { false }; bool childDone[8] = { false };
        float cost[8][8];
        for (size_t i = 0; i < ch.size(); i++) {
            float d[3];
            for (int k = 0; k < 3; k++) d[k] = (ch[i].box.lo[k] + ch[i].box.hi[k]) - (nb.lo[k] + nb.hi[k]);
            for (int s = 0; s < 8; s++)
                cost[i][s] = ((s & 4) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 1) ? d[2] : -d[2]);
        }
        // nodes with at most four children use slots 0..3 only: the kernel skips the upper half of such a node
        const int numSlots = ch.size() <= 4 ? 4 : 8;
        for (size_t round = 0; round < ch.size(); round++) {
            int bi = -1, bs = -1; float bc = -3.0e38f;
            for (size_t i = 0; i < ch.size(); i++) if (!childDone[i])
                for (int s = 0; s < numSlots; s++) if (!slotUsed[s] && cost[i][s] > bc) { bc = cost[i][s]; bi = (int)i; bs = s; }
            slotOf[bi] = bs; slotUsed[bs] = true; childDone[bi] = true;
        }
        int childAt[8]; for (int s = 0; s < 8; s++) childAt[s] = -1;
        for (size_t i = 0; i < ch.size(); i++) childAt[slotOf[i]] = (int)i;

        // ---- emit ----
        uint32_t w[20]; memset(w, 0, sizeof(w));
        memcpy(&w[0], &nb.lo[0], 4); memcpy(&w[1], &nb.lo[1], 4); memcpy(&w[2], &nb.lo[2], 4);
        uint8_t meta[8] = { 0 }, q[6][8]; memset(q, 0, sizeof(q));
        uint32_t imask = 0;
        const uint32_t childBase = (uint32_t)(out.nodes.size() / 20);
        const uint32_t triBase = (uint32_t)out.triIndex.size();
        int triOffset = 0;
        for (int s = 0; s < 8; s++) {
            if (childAt[s] < 0) continue;
            const Child& c = ch[childAt[s]];
            for (int k = 0; k < 3; k++) {
                const double step = ldexp(1.0, e[k]);
                double lo = floor(((double)c.box.lo[k] - (double)nb.lo[k]) / step);
                double hi = ceil(((double)c.box.hi[k] - (double)nb.lo[k]) / step);
                q[k][s] = (uint8_t)std::max(0.0, std::min(255.0, lo));
                q[3 + k][s] = (uint8_t)std::max(0.0, std::min(255.0, hi));
            }
            if (c.inner >= 0) {
                imask |= 1u << s;
                meta[s] = (uint8_t)(0x20 | 24 | s);                 // 0b001_11sss
                out.nodes.resize(out.nodes.size() + 20, 0u);        // children are written in slot order
                queue.push_back(Pending{ c.inner, cur.depth + 1 });
            } else {
                const uint8_t unary = (uint8_t)((1u << c.count) - 1u);      // 1, 3, 7
                meta[s] = (uint8_t)((unary << 5) | triOffset);
                for (int t = 0; t < c.count; t++) {
                    const size_t row = (size_t)c.rowStart + 3 * (size_t)t;
                    out.woop.insert(out.woop.end(), woop + row * 4, woop + row * 4 + 12);
                    out.triIndex.push_back(triIndex[row]);
                }
                triOffset += c.count;
            }
        }
        w[3] = (uint32_t)(e[0] + 142) | ((uint32_t)(e[1] + 142) << 8) | ((uint32_t)(e[2] + 142) << 16) | (imask << 24);
        w[4] = childBase; w[5] = triBase;
        memcpy(&w[6], meta, 8);
        memcpy(&w[8], q[0], 8);  memcpy(&w[10], q[1], 8);       // qlo_x, qlo_y
        memcpy(&w[12], q[2], 8); memcpy(&w[14], q[3], 8);       // qlo_z, qhi_x
        memcpy(&w[16], q[4], 8); memcpy(&w[18], q[5], 8);       // qhi_y, qhi_z
        memcpy(&out.nodes[next * 20], w, sizeof(w));
        next++;
    }
    out.numNodes = (int)(out.nodes.size() / 20);
    out.numTris = (int)out.triIndex.size();
    out.avgChildren = out.numNodes ? childSum / out.numNodes : 0.0;
    // two rows of padding, like the binary path (a lane may prefetch past the last triangle)
    out.woop.insert(out.woop.end(), 8, 0.0f);
    return true;
}

}  // namespace b2rt

namespace b2rt {

bool validateWide8(const float* nodes2, size_t numNodes2, const float* woop, const int32_t* triIndex, size_t numRows,
                   const Wide8& w, const char** error)
{
    Binary bin{ nodes2, numNodes2, woop, numRows };
    // triangle references of the binary tree, in any order
    std::vector<int32_t> want;
    std::vector<int> stack(1, 0);
    while (!stack.empty()) {
        const int node = stack.back(); stack.pop_back();
        for (int c = 0; c < 2; c++) {
            const int ref = bin.childRef(node, c);
            if (ref >= 0) { stack.push_back(ref / 4); continue; }
            const int n = bin.leafCount(~ref);
            for (int t = 0; t < n; t++) want.push_back(triIndex[(size_t)(~ref) + 3 * (size_t)t]);
        }
    }
    std::vector<int32_t> got(w.triIndex);
    std::sort(want.begin(), want.end());
    std::sort(got.begin(), got.end());
    if (want != got) { if (error) *error = "wide BVH does not reference the same triangles as the binary tree"; return false; }
    // indices in range, every node reachable exactly once
    std::vector<char> seen((size_t)w.numNodes, 0);
    std::vector<uint32_t> st(1, 0u);
    size_t reached = 0;
    while (!st.empty()) {
        const uint32_t n = st.back(); st.pop_back();
        if (n >= (uint32_t)w.numNodes || seen[n]) { if (error) *error = "wide BVH child index out of range or shared"; return false; }
        seen[n] = 1; reached++;
        const uint32_t* p = &w.nodes[(size_t)n * 20];
        const uint32_t imask = p[3] >> 24, childBase = p[4], triBase = p[5];
        uint8_t meta[8]; memcpy(meta, p + 6, 8);
        uint32_t rel = 0;
        for (int s = 0; s < 8; s++) {
            if (imask & (1u << s)) { if (meta[s] != (0x38 | s)) { if (error) *error = "inner meta byte"; return false; } st.push_back(childBase + rel++); }
            else if (meta[s]) {
                const uint32_t cnt = (meta[s] >> 5) == 1 ? 1 : ((meta[s] >> 5) == 3 ? 2 : 3);
                if (triBase + (meta[s] & 31) + cnt > (uint32_t)w.numTris) { if (error) *error = "triangle range"; return false; }
            }
        }
    }
    if (reached != (size_t)w.numNodes) { if (error) *error = "unreachable wide nodes"; return false; }
    return true;
}

}  // namespace b2rt

// ---- C ABI (host only): build / inspect the wide form without a GPU (tests, tooling) ----
#include "../../../include/b2rt.h"

struct b2rt_wide_bvh { b2rt::Wide8 w; };

extern "C" {

int b2rt_wide_build(const void* nodes, size_t node_bytes, const void* woop, size_t woop_bytes, const int32_t* tri_index,
                    size_t tri_index_bytes, b2rt_wide_bvh** out)
{
    if (!out) return B2RT_ERR_ARG;
    *out = nullptr;
    if (!nodes || !woop || !tri_index || node_bytes < 64 || node_bytes % 64 || woop_bytes % 16 || tri_index_bytes * 4 != woop_bytes)
        return B2RT_ERR_ARG;
    b2rt_wide_bvh* h = new b2rt_wide_bvh();
    const char* err = nullptr;
    const float* n = static_cast<const float*>(nodes);
    const float* w = static_cast<const float*>(woop);
    if (!b2rt::buildWide8(n, node_bytes / 64, w, tri_index, woop_bytes / 16, h->w, &err) ||
        !b2rt::validateWide8(n, node_bytes / 64, w, tri_index, woop_bytes / 16, h->w, &err)) {
        delete h;
        return B2RT_ERR_ARG;
    }
    *out = h;
    return B2RT_OK;
}

void b2rt_wide_free(b2rt_wide_bvh* h) { delete h; }

int b2rt_wide_arrays(const b2rt_wide_bvh* h, const void** nodes, size_t* node_bytes, const void** woop, size_t* woop_bytes,
                     const int32_t** tri_index, size_t* tri_index_bytes, int stats[3])
{
    if (!h) return B2RT_ERR_ARG;
    if (nodes) *nodes = h->w.nodes.data();
    if (node_bytes) *node_bytes = h->w.nodes.size() * 4;
    if (woop) *woop = h->w.woop.data();
    if (woop_bytes) *woop_bytes = h->w.woop.size() * 4;
    if (tri_index) *tri_index = h->w.triIndex.data();
    if (tri_index_bytes) *tri_index_bytes = h->w.triIndex.size() * 4;
    if (stats) { stats[0] = h->w.numNodes; stats[1] = h->w.numTris; stats[2] = h->w.maxDepth; }
    return B2RT_OK;
}

}  // extern "C"
