// wide_bvh.cpp -- host-side collapse of a Compact2 BVH into the private 8-wide compressed form
// described in ../wide_bvh.h.  Pure host C++ (no CUDA); one-off work inside b2rt_set_bvh.
#include "../wide_bvh.h"

#include <math.h>
#include <string.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <deque>

namespace b2rt {

namespace {

struct Box { float lo[3], hi[3]; };

constexpr double kMargin = 1.0 / 64.0;      // slack of every quantised plane, in cells (see buildWide8)

inline float area(const Box& b)
{
    const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    return dx * dy + dy * dz + dz * dx;
}

// One child of a wide node while it is being assembled.
struct Child {
    Box box;
    int inner;          // >= 0: binary node index to collapse further; -1: leaf part
    int rows[3];        // leaf part: first Woop row of each of its triangles (caller's array)
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
            ch.inner = ref / 4; ch.rows[0] = ch.rows[1] = ch.rows[2] = 0; ch.count = 0;
            out.push_back(ch);
        } else {
            const int row = ~ref;
            const int n = bin.leafCount(row);
            if (n < 0) return false;
            for (int first = 0; first < n; first += 3) {
                ch.inner = -1; ch.count = std::min(3, n - first);
                for (int t = 0; t < ch.count; t++) ch.rows[t] = row + (first + t) * 3;
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


// ---- optimal collapse (after Ylitie, Karras, Laine 2017, section 3.1) ----------------------------------------------
// cost(n, i) = cheapest SAH cost of the binary subtree under n when it may occupy at most i child slots of the wide
// node above it.  One slot holds either an inner wide node (cost: its box area, plus its own subtree distributed over
// eight slots) or up to three triangles (cost: area x triangles x kPrimCost).  A binary leaf of the caller's tree has
// 1..8 triangles and therefore needs 1..3 slots with the same box; a binary subtree with at most three triangle
// references in total may be merged into ONE leaf slot.  The greedy top-down collapse this replaces ("open the largest
// child while the result fits") filled a node with 4.3 children on average on the 910 K-triangle knot -- two thirds of the
// nodes had four or fewer -- because whatever hangs below the eighth slot becomes a node of its own however small it is.
constexpr float kNodeCost = 1.0f, kPrimCost = 0.3f;
constexpr float kInf = 3.0e38f;

struct Collapse {
    const Binary& bin;
    int width;
    std::vector<float> cost;            // [node][i-1], i = 1..7
    std::vector<uint8_t> split;         // [node][i-1]: i >= 2: slots given to child 0 (0 = "same as i-1"); i == 1: 1 = merged leaf, 0 = inner node
    std::vector<uint8_t> split8;        // [node]: slots given to child 0 when the node becomes a wide node of its own
    std::vector<float> nodeArea;        // area of the node's own box (from its parent's record; root: union of its children)
    std::vector<int> refs;              // triangle references below the node (saturates at 255)

    explicit Collapse(const Binary& b, int w) : bin(b), width(w) {}

    // cost of child c of node n with at most i slots (i >= 1)
    float childCost(int n, int c, int i) const
    {
        const int ref = bin.childRef(n, c);
        if (i > 7) i = 7;
        if (ref >= 0) return cost[(size_t)(ref / 4) * 7 + (i - 1)];
        const int cnt = bin.leafCount(~ref);
        if (cnt == 0) return 0.0f;                      // a leaf without triangles takes no slot
        return (i >= (cnt + 2) / 3) ? area(bin.childBox(n, c)) * (float)cnt * kPrimCost : kInf;
    }
    int minSlots(int n, int c) const
    {
        const int ref = bin.childRef(n, c);
        return ref >= 0 ? 1 : (std::max(0, bin.leafCount(~ref)) + 2) / 3;
    }
    int childRefs(int n, int c) const
    {
        const int ref = bin.childRef(n, c);
        return ref >= 0 ? refs[ref / 4] : std::max(0, bin.leafCount(~ref));
    }

    bool run()
    {
        const size_t n = bin.numNodes;
        cost.assign(n * 7, kInf); split.assign(n * 7, 0); split8.assign(n, 1); nodeArea.assign(n, 0.0f); refs.assign(n, 0);
        // post-order over the binary tree (iterative; the tree may be 64 levels deep and millions of nodes large)
        std::vector<int> order; order.reserve(n);
        std::vector<char> seen(n, 0);
        std::vector<int> stack(1, 0);
        {
            Box a = bin.childBox(0, 0), b = bin.childBox(0, 1), u;
            for (int k = 0; k < 3; k++) { u.lo[k] = std::min(a.lo[k], b.lo[k]); u.hi[k] = std::max(a.hi[k], b.hi[k]); }
            nodeArea[0] = area(u);
        }
        while (!stack.empty()) {
            const int node = stack.back(); stack.pop_back();
            if (seen[node]) return false;               // not a tree
            seen[node] = 1; order.push_back(node);
            for (int c = 0; c < 2; c++) {
                const int ref = bin.childRef(node, c);
                if (ref < 0) { if (bin.leafCount(~ref) < 0) return false; continue; }
                if (ref % 4 != 0 || (size_t)(ref / 4) >= n) return false;
                nodeArea[ref / 4] = area(bin.childBox(node, c));
                stack.push_back(ref / 4);
            }
        }
        for (size_t k = order.size(); k-- > 0;) {
            const int node = order[k];
            refs[node] = std::min(255, childRefs(node, 0) + childRefs(node, 1));
            float* cn = &cost[(size_t)node * 7];
            uint8_t* sp = &split[(size_t)node * 7];
            // distribute(j): children 0 and 1 share j slots; a node of its own has `own` of them
            // (two fat leaves side by side need up to six whatever the width asked for)
            const int own = std::max(width, minSlots(node, 0) + minSlots(node, 1));
            float dist[9]; uint8_t distK[9];
            for (int j = 2; j <= own; j++) {
                dist[j] = kInf; distK[j] = 1;
                for (int a = 0; a <= j; a++) {       // a child without triangles takes no slot
                    const float ca = (a == 0) ? (childRefs(node, 0) == 0 ? 0.0f : kInf) : childCost(node, 0, a);
                    const float cb = (j - a == 0) ? (childRefs(node, 1) == 0 ? 0.0f : kInf) : childCost(node, 1, j - a);
                    if (ca < kInf && cb < kInf && ca + cb < dist[j]) { dist[j] = ca + cb; distK[j] = (uint8_t)a; }
                }
            }
            for (int j = own + 1; j <= 8; j++) dist[j] = kInf;
            split8[node] = (uint8_t)(distK[own] | ((own - distK[own]) << 4));
            if (node == 0) rootOk = dist[own] < kInf;
            dist[8] = dist[own];
            const float internal = dist[8] < kInf ? dist[8] + nodeArea[node] * kNodeCost : kInf;
            const float merged = (refs[node] >= 1 && refs[node] <= 3) ? nodeArea[node] * (float)refs[node] * kPrimCost : kInf;
            if (merged < internal) { cn[0] = merged; sp[0] = 1; } else { cn[0] = internal; sp[0] = 0; }
            for (int i = 2; i <= 7; i++) {
                if (i <= own && dist[i] < cn[i - 2]) { cn[i - 1] = dist[i]; sp[i - 1] = distK[i] ? distK[i] : (uint8_t)0x80; }
                else { cn[i - 1] = cn[i - 2]; sp[i - 1] = 0; }
            }
        }
        return rootOk;
    }
    bool rootOk = false;

    // the triangles below child c of node n, appended as rows
    void gatherRows(int n, int c, std::vector<int>& rows) const
    {
        const int ref = bin.childRef(n, c);
        if (ref < 0) { const int cnt = bin.leafCount(~ref); for (int t = 0; t < cnt; t++) rows.push_back(~ref + 3 * t); return; }
        gatherRows(ref / 4, 0, rows); gatherRows(ref / 4, 1, rows);      // at most three references below: shallow
    }

    // children of the wide node that stands for binary node `node`
    void children(int node, std::vector<Child>& out) const
    {
        expand(node, 0, split8[node] & 15, out);
        expand(node, 1, split8[node] >> 4, out);
    }
    // child c of binary node n, with at most i slots
    void expand(int n, int c, int i, std::vector<Child>& out) const
    {
        if (i <= 0) return;
        if (i > 7) i = 7;
        const int ref = bin.childRef(n, c);
        Child ch;
        ch.box = bin.childBox(n, c);
        if (ref < 0) {
            const int row = ~ref, cnt = bin.leafCount(row);
            for (int first = 0; first < cnt; first += 3) {
                ch.inner = -1; ch.count = std::min(3, cnt - first);
                for (int t = 0; t < ch.count; t++) ch.rows[t] = row + (first + t) * 3;
                out.push_back(ch);
            }
            return;
        }
        const int m = ref / 4;
        while (i > 1 && split[(size_t)m * 7 + (i - 1)] == 0) i--;          // "no better than with one slot fewer"
        if (i == 1) {
            if (split[(size_t)m * 7] == 1) {                                // merged leaf
                std::vector<int> rows;
                gatherRows(m, 0, rows); gatherRows(m, 1, rows);
                ch.inner = -1; ch.count = (int)rows.size();
                for (int t = 0; t < ch.count && t < 3; t++) ch.rows[t] = rows[t];
                out.push_back(ch);
            } else {
                ch.inner = m; ch.rows[0] = ch.rows[1] = ch.rows[2] = 0; ch.count = 0;
                out.push_back(ch);
            }
            return;
        }
        const int a = split[(size_t)m * 7 + (i - 1)] & 0x7F;
        expand(m, 0, a, out);
        expand(m, 1, i - a, out);
    }
};

}  // namespace

bool buildWide8(const float* nodes2, size_t numNodes2, const float* woop, const int32_t* triIndex, size_t numRows,
                Wide8& out, const char** error, bool greedy)
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
    // greedy = the round-1 top-down collapse, kept for A/B measurements (B2RT_TUNE_GREEDY_WIDE)
    const int width = 8;
    Collapse dp(bin, width);
    if (!greedy && !dp.run()) { if (error) *error = kBad; return false; }
    while (!queue.empty()) {
        const Pending cur = queue.front(); queue.pop_front();
        ch.clear();
        if (!greedy) {
            dp.children(cur.binaryNode, ch);
        } else {
        // ---- greedy collapse: open the inner child with the largest box while the result fits eight slots ----
        if (!appendChildren(bin, cur.binaryNode, ch)) { if (error) *error = kBad; return false; }
        while (true) {
            int best = -1; float bestArea = -1.0f;
            for (size_t i = 0; i < ch.size(); i++) {
                if (ch[i].inner < 0) continue;
                const int need = slotsOfBinaryChild(bin, ch[i].inner, 0) + slotsOfBinaryChild(bin, ch[i].inner, 1);
                if ((int)ch.size() - 1 + need > width) continue;
                const float a = area(ch[i].box);
                if (a > bestArea) { bestArea = a; best = (int)i; }
            }
            if (best < 0) break;
            const int open = ch[best].inner;
            ch.erase(ch.begin() + best);
            if (!appendChildren(bin, open, ch)) { if (error) *error = kBad; return false; }
        }
        }
        if (ch.size() > 8) { if (error) *error = "internal: wide node overflow"; return false; }
        childSum += (double)ch.size();
        if (cur.depth > out.maxDepth) out.maxDepth = cur.depth;

        // ---- node frame: origin and per-axis power-of-two grid step ----
        Box nb; for (int k = 0; k < 3; k++) { nb.lo[k] = 3.0e38f; nb.hi[k] = -3.0e38f; }
        for (const Child& c : ch) for (int k = 0; k < 3; k++) { nb.lo[k] = std::min(nb.lo[k], c.box.lo[k]); nb.hi[k] = std::max(nb.hi[k], c.box.hi[k]); }
        if (ch.empty()) for (int k = 0; k < 3; k++) nb.lo[k] = nb.hi[k] = 0.0f;
        // Quantisation frame.  The kernel decodes a plane as t = (1 + q 2^-15) * (2^(e+15) idir) + ((p - orig) idir - 2^(e+15) idir)
        // (trace_wide.cu): the second term is ~2^15 cells large, so its rounding moves a decoded plane by up to 2^-9 of a cell in
        // either direction.  Every child box is therefore quantised with kMargin = 2^-6 of a cell of slack on both sides
        // (floor(x - margin), ceil(x + margin)) and the node origin sits two margins below the children's lower bound, so the
        // DECODED box -- kernel arithmetic included -- always contains the binary tree's box (checked by validateWide8).
        int e[3];
        float origin[3];
        for (int k = 0; k < 3; k++) {
            const double ext = (double)nb.hi[k] - (double)nb.lo[k];
            if (!(ext > 0.0)) { e[k] = -141; }
            else {
                int ek = (int)ceil(log2(ext / 254.0));
                while (ceil(ext / ldexp(1.0, ek) + 4.0 * kMargin) > 255.0) ek++;
                e[k] = std::max(-141, std::min(112, ek));
            }
            // origin: at least two margins below the lowest child plane, in float
            auto place = [&](int ek) {
                const double st = ldexp(1.0, ek);
                float o = (float)((double)nb.lo[k] - 2.0 * kMargin * st);
                while ((double)nb.lo[k] - (double)o < 2.0 * kMargin * st) o = nextafterf(o, -INFINITY);
                return o;
            };
            float o = place(e[k]);
            // a box so far from zero that the margin vanishes in float (or a flat one): widen the step until the top plane fits
            while ((((double)nb.hi[k] - (double)o) / ldexp(1.0, e[k]) + kMargin) > 255.0 && e[k] < 112) o = place(++e[k]);
            origin[k] = o;
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
        memcpy(&w[0], &origin[0], 4); memcpy(&w[1], &origin[1], 4); memcpy(&w[2], &origin[2], 4);
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
                double lo = floor(((double)c.box.lo[k] - (double)origin[k]) / step - kMargin);
                double hi = ceil(((double)c.box.hi[k] - (double)origin[k]) / step + kMargin);
                if (lo < 0.0 || hi > 255.0) { if (error) *error = "internal: quantised child box leaves the node grid"; return false; }
                q[k][s] = (uint8_t)lo;
                q[3 + k][s] = (uint8_t)hi;
                // the decoded planes keep at least half a margin of slack against the binary tree's box
                if ((double)origin[k] + lo * step > (double)c.box.lo[k] - 0.5 * kMargin * step ||
                    (double)origin[k] + hi * step < (double)c.box.hi[k] + 0.5 * kMargin * step) {
                    if (error) *error = "internal: quantised child box is not a conservative superset";
                    return false;
                }
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
                    const size_t row = (size_t)c.rows[t];
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
#include "../../../include/b2rt_tools.h"

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
    const bool greedy = getenv("B2RT_WIDE_GREEDY") != nullptr;      // tooling: the round-1 collapse, for A/B counts on the CPU
    if (!b2rt::buildWide8(n, node_bytes / 64, w, tri_index, woop_bytes / 16, h->w, &err, greedy) ||
        !b2rt::validateWide8(n, node_bytes / 64, w, tri_index, woop_bytes / 16, h->w, &err)) {
        if (getenv("B2RT_DEBUG")) fprintf(stderr, "b2rt_wide_build: %s\n", err ? err : "?");
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
