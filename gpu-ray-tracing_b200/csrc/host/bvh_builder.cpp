// bvh_builder.cpp -- host-side SBVH build and flatten to the Compact2 device layout.
//
// Produces what the reference's SplitBVHBuilder (rt/bvh/SplitBVHBuilder.cc) + CudaBVH::createCompact
// / woopifyTri (rt/cuda/CudaBVH.cc:270-380) produce: 64-byte child-pair nodes, 48-byte Woop
// triangles inlined per leaf with a 0x80000000 terminator row, and a triIndex array that mirrors
// the Woop rows.  The layout is the contract with the traversal kernel; the tree is this
// builder's own (same algorithm family -- SAH object splits, binned spatial splits with
// reference unsplitting, Stich et al. 2009 -- same cost model: Platform("GPU"), node cost 1,
// triangle cost 1, leaves of 1..8, Renderer.cc:53-54), built for speed:
//   * binned SAH (64 bins) for large nodes, exact sweep for small ones (the reference sorts all
//     references three times per node, SplitBVHBuilder.cc:212-214);
//   * subtrees are built by a pool of std::threads (the reference is single threaded);
//   * nodes are emitted depth first with the first child adjacent to its parent.
// Compiled with -ffp-contract=off: the Woop transform follows the reference's cofactor
// expansion operation for operation (Math.hh:962-983) so the rows are bit-identical to
// CudaBVH::woopifyTri's.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include "../../../include/b2rt.h"

namespace {

constexpr float kBig = 3.402823466e+38f;
constexpr int kMaxDepth = 64;           // SplitBVHBuilder.hh:43
constexpr int kMaxSpatialDepth = 48;    // SplitBVHBuilder.hh:44
constexpr int kSpatialBins = 128;       // SplitBVHBuilder.hh:45
constexpr int kObjectBins = 64;
constexpr int kSweepThreshold = 2048;   // nodes up to this many references use the exact sweep
constexpr int kParallelThreshold = 16384;

struct Box {
    float lo[3], hi[3];
    Box() { for (int k = 0; k < 3; k++) { lo[k] = kBig; hi[k] = -kBig; } }
    void grow(const float* p) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } }
    void grow(const Box& b) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
    void clip(const Box& b) { for (int k = 0; k < 3; k++) { lo[k] = std::max(lo[k], b.lo[k]); hi[k] = std::min(hi[k], b.hi[k]); } }
    bool valid() const { return lo[0] <= hi[0] && lo[1] <= hi[1] && lo[2] <= hi[2]; }
    float area() const
    {
        if (!valid()) return 0.0f;
        const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return (dx * dy + dy * dz + dz * dx) * 2.0f;
    }
};

struct Ref { int tri; Box box; };

struct Node {
    Box box;
    Node* child[2] = { nullptr, nullptr };
    int* tris = nullptr;        // leaf payload
    int numTris = 0;
    bool leaf() const { return child[0] == nullptr; }
};

// per-thread bump allocators; they live until the tree has been flattened
struct Arena {
    std::vector<std::unique_ptr<Node[]>> nodeChunks;
    std::vector<std::unique_ptr<int[]>> intChunks;
    size_t nodeUsed = 0, nodeCap = 0, intUsed = 0, intCap = 0;
    Node* newNode()
    {
        if (nodeUsed == nodeCap) { nodeCap = 1 << 14; nodeUsed = 0; nodeChunks.emplace_back(new Node[nodeCap]); }
        return &nodeChunks.back()[nodeUsed++];
    }
    int* newInts(int n)
    {
        if (intUsed + (size_t)n > intCap) { intCap = std::max<size_t>(1 << 16, (size_t)n); intUsed = 0; intChunks.emplace_back(new int[intCap]); }
        int* p = &intChunks.back()[intUsed];
        intUsed += (size_t)n;
        return p;
    }
};

struct Builder {
    const float* verts; const int32_t* tris; int numTris;
    float alpha; int minLeaf, maxLeaf; bool spatial; int maxThreads;
    float minOverlap = 0.0f;
    std::atomic<int> activeThreads{ 1 };
    std::mutex arenaMutex;
    std::vector<std::unique_ptr<Arena>> arenas;
    std::atomic<long long> duplicates{ 0 };

    Arena* newArena()
    {
        std::lock_guard<std::mutex> g(arenaMutex);
        arenas.emplace_back(new Arena());
        return arenas.back().get();
    }
    const float* vtx(int tri, int k) const { return verts + 3 * (size_t)tris[3 * (size_t)tri + k]; }

    // Bounds of the parts of `ref` on either side of the plane x[dim] = pos
    // (SplitBVHBuilder::splitReference, SplitBVHBuilder.cc:441-486).
    void splitRef(const Ref& ref, int dim, float pos, Ref& left, Ref& right) const
    {
        left.tri = right.tri = ref.tri;
        left.box = Box(); right.box = Box();
        const float* prev = vtx(ref.tri, 2);
        for (int i = 0; i < 3; i++) {
            const float* cur = vtx(ref.tri, i);
            const float a = prev[dim], b = cur[dim];
            if (a <= pos) left.box.grow(prev);
            if (a >= pos) right.box.grow(prev);
            if ((a < pos && b > pos) || (a > pos && b < pos)) {
                float t = (pos - a) / (b - a);
                t = std::min(std::max(t, 0.0f), 1.0f);
                const float p[3] = { prev[0] + (cur[0] - prev[0]) * t, prev[1] + (cur[1] - prev[1]) * t, prev[2] + (cur[2] - prev[2]) * t };
                left.box.grow(p); right.box.grow(p);
            }
            prev = cur;
        }
        left.box.hi[dim] = pos;
        right.box.lo[dim] = pos;
        left.box.clip(ref.box);
        right.box.clip(ref.box);
    }

    Node* makeLeaf(Arena* ar, const std::vector<Ref>& refs, const Box& box)
    {
        Node* n = ar->newNode();
        n->box = box;
        n->numTris = (int)refs.size();
        n->tris = ar->newInts(std::max(1, n->numTris));
        for (int i = 0; i < n->numTris; i++) n->tris[i] = refs[i].tri;
        return n;
    }

    struct ObjectSplit { float sah = kBig; int dim = 0; int numLeft = 0; float pos = 0; bool binned = false; Box lb, rb; };
    struct SpatialSplit { float sah = kBig; int dim = 0; float pos = 0; };

    static float centroid2(const Ref& r, int d) { return r.box.lo[d] + r.box.hi[d]; }

    ObjectSplit findObjectSplitSweep(std::vector<Ref>& refs, float nodeSAH, std::vector<Box>& rightBoxes) const
    {
        ObjectSplit best;
        float bestTie = kBig;
        const int n = (int)refs.size();
        rightBoxes.resize(n);
        for (int dim = 0; dim < 3; dim++) {
            std::sort(refs.begin(), refs.end(), [dim](const Ref& a, const Ref& b) {
                const float ca = centroid2(a, dim), cb = centroid2(b, dim);
                return ca < cb || (ca == cb && a.tri < b.tri);
            });
            Box rb;
            for (int i = n - 1; i > 0; i--) { rb.grow(refs[i].box); rightBoxes[i - 1] = rb; }
            Box lb;
            for (int i = 1; i < n; i++) {
                lb.grow(refs[i - 1].box);
                const float sah = nodeSAH + lb.area() * (float)i + rightBoxes[i - 1].area() * (float)(n - i);
                const float tie = (float)i * (float)i + (float)(n - i) * (float)(n - i);
                if (sah < best.sah || (sah == best.sah && tie < bestTie)) {
                    best.sah = sah; best.dim = dim; best.numLeft = i; best.lb = lb; best.rb = rightBoxes[i - 1];
                    bestTie = tie;
                }
            }
        }
        return best;
    }

    ObjectSplit findObjectSplitBinned(const std::vector<Ref>& refs, float nodeSAH) const
    {
        ObjectSplit best;
        best.binned = true;
        const int n = (int)refs.size();
        Box cb;     // centroid bounds (of lo+hi)
        for (const Ref& r : refs) { const float c[3] = { centroid2(r, 0), centroid2(r, 1), centroid2(r, 2) }; cb.grow(c); }
        for (int dim = 0; dim < 3; dim++) {
            const float ext = cb.hi[dim] - cb.lo[dim];
            if (!(ext > 0.0f)) continue;
            const float scale = (float)kObjectBins / ext;
            Box bins[kObjectBins]; int counts[kObjectBins] = { 0 };
            for (const Ref& r : refs) {
                int b = (int)((centroid2(r, dim) - cb.lo[dim]) * scale);
                b = std::min(std::max(b, 0), kObjectBins - 1);
                bins[b].grow(r.box); counts[b]++;
            }
            Box rbs[kObjectBins]; Box rb;
            for (int i = kObjectBins - 1; i > 0; i--) { rb.grow(bins[i]); rbs[i - 1] = rb; }
            Box lb; int nl = 0;
            for (int i = 1; i < kObjectBins; i++) {
                lb.grow(bins[i - 1]); nl += counts[i - 1];
                if (nl == 0 || nl == n) continue;
                const float sah = nodeSAH + lb.area() * (float)nl + rbs[i - 1].area() * (float)(n - nl);
                if (sah < best.sah) {
                    best.sah = sah; best.dim = dim; best.numLeft = nl; best.lb = lb; best.rb = rbs[i - 1];
                    best.pos = cb.lo[dim] + ext * ((float)i / (float)kObjectBins);
                }
            }
        }
        return best;
    }

    SpatialSplit findSpatialSplit(const std::vector<Ref>& refs, const Box& box, float nodeSAH) const
    {
        struct Bin { Box box; int enter = 0, exit = 0; };
        std::vector<Bin> bins(3 * kSpatialBins);
        float binSize[3], invBin[3];
        for (int k = 0; k < 3; k++) { binSize[k] = (box.hi[k] - box.lo[k]) * (1.0f / (float)kSpatialBins); invBin[k] = 1.0f / binSize[k]; }
        for (const Ref& ref : refs) {
            for (int dim = 0; dim < 3; dim++) {
                if (!(binSize[dim] > 0.0f)) continue;
                int first = (int)((ref.box.lo[dim] - box.lo[dim]) * invBin[dim]);
                int last = (int)((ref.box.hi[dim] - box.lo[dim]) * invBin[dim]);
                first = std::min(std::max(first, 0), kSpatialBins - 1);
                last = std::min(std::max(last, first), kSpatialBins - 1);
                Ref cur = ref;
                for (int i = first; i < last; i++) {
                    Ref l, r;
                    splitRef(cur, dim, box.lo[dim] + binSize[dim] * (float)(i + 1), l, r);
                    bins[dim * kSpatialBins + i].box.grow(l.box);
                    cur = r;
                }
                bins[dim * kSpatialBins + last].box.grow(cur.box);
                bins[dim * kSpatialBins + first].enter++;
                bins[dim * kSpatialBins + last].exit++;
            }
        }
        SpatialSplit best;
        Box rbs[kSpatialBins];
        const int n = (int)refs.size();
        for (int dim = 0; dim < 3; dim++) {
            if (!(binSize[dim] > 0.0f)) continue;
            const Bin* b = &bins[dim * kSpatialBins];
            Box rb;
            for (int i = kSpatialBins - 1; i > 0; i--) { rb.grow(b[i].box); rbs[i - 1] = rb; }
            Box lb; int nl = 0, nr = n;
            for (int i = 1; i < kSpatialBins; i++) {
                lb.grow(b[i - 1].box);
                nl += b[i - 1].enter; nr -= b[i - 1].exit;
                const float sah = nodeSAH + lb.area() * (float)nl + rbs[i - 1].area() * (float)nr;
                if (sah < best.sah) { best.sah = sah; best.dim = dim; best.pos = box.lo[dim] + binSize[dim] * (float)i; }
            }
        }
        return best;
    }

    // SplitBVHBuilder::performSpatialSplit (SplitBVHBuilder.cc:345-437): straddling references
    // are duplicated, or moved whole to one side when that is cheaper ("unsplitting").
    void doSpatialSplit(std::vector<Ref>& refs, const SpatialSplit& sp, std::vector<Ref>& L, std::vector<Ref>& R, Box& lb, Box& rb)
    {
        std::vector<Ref> straddle;
        lb = Box(); rb = Box();
        for (const Ref& r : refs) {
            if (r.box.hi[sp.dim] <= sp.pos) { lb.grow(r.box); L.push_back(r); }
            else if (r.box.lo[sp.dim] >= sp.pos) { rb.grow(r.box); R.push_back(r); }
            else straddle.push_back(r);
        }
        for (const Ref& r : straddle) {
            Ref l, rr;
            splitRef(r, sp.dim, sp.pos, l, rr);
            Box lub = lb, rub = rb, ldb = lb, rdb = rb;
            lub.grow(r.box); rub.grow(r.box); ldb.grow(l.box); rdb.grow(rr.box);
            const float lac = (float)L.size(), rac = (float)R.size(), lbc = (float)(L.size() + 1), rbc = (float)(R.size() + 1);
            const float unsplitL = lub.area() * lbc + rb.area() * rac;
            const float unsplitR = lb.area() * lac + rub.area() * rbc;
            const float dup = ldb.area() * lbc + rdb.area() * rbc;
            const float m = std::min(unsplitL, std::min(unsplitR, dup));
            if (m == unsplitL) { lb = lub; L.push_back(r); }
            else if (m == unsplitR) { rb = rub; R.push_back(r); }
            else { lb = ldb; rb = rdb; L.push_back(l); R.push_back(rr); }
        }
    }

    Node* build(Arena* ar, std::vector<Ref>& refs, Box box, int depth)
    {
        // drop references whose box has degenerated to a segment or a point (SplitBVHBuilder.cc:136-146)
        {
            size_t w = 0;
            for (size_t i = 0; i < refs.size(); i++) {
                const Box& b = refs[i].box;
                const float sx = b.hi[0] - b.lo[0], sy = b.hi[1] - b.lo[1], sz = b.hi[2] - b.lo[2];
                const float mn = std::min(sx, std::min(sy, sz)), mx = std::max(sx, std::max(sy, sz));
                if (mn < 0.0f || (sx + sy + sz) == mx) continue;
                if (w != i) refs[w] = refs[i];
                w++;
            }
            refs.resize(w);
        }
        const int n = (int)refs.size();
        if (n <= minLeaf || depth >= kMaxDepth) return makeLeaf(ar, refs, box);

        const float area = box.area();
        const float leafSAH = area * (float)n;
        const float nodeSAH = area * 2.0f;
        std::vector<Box> tmpBoxes;
        ObjectSplit obj = (n <= kSweepThreshold) ? findObjectSplitSweep(refs, nodeSAH, tmpBoxes) : findObjectSplitBinned(refs, nodeSAH);

        SpatialSplit spl;
        if (spatial && depth < kMaxSpatialDepth && obj.sah < kBig) {
            Box overlap = obj.lb;
            overlap.clip(obj.rb);
            if (overlap.area() >= minOverlap) spl = findSpatialSplit(refs, box, nodeSAH);
        }

        const float minSAH = std::min(leafSAH, std::min(obj.sah, spl.sah));
        if (minSAH == leafSAH && n <= maxLeaf) return makeLeaf(ar, refs, box);

        std::vector<Ref> L, R;
        Box lb, rb;
        if (minSAH == spl.sah && spl.sah < kBig) {
            doSpatialSplit(refs, spl, L, R, lb, rb);
            if (L.empty() || R.empty()) { L.clear(); R.clear(); }
        }
        if (L.empty() || R.empty()) {
            if (obj.sah < kBig && !obj.binned) {
                // the sweep left `refs` sorted along its last dimension; re-sort along the winner
                const int dim = obj.dim;
                std::sort(refs.begin(), refs.end(), [dim](const Ref& a, const Ref& b) {
                    const float ca = centroid2(a, dim), cb = centroid2(b, dim);
                    return ca < cb || (ca == cb && a.tri < b.tri);
                });
                L.assign(refs.begin(), refs.begin() + obj.numLeft);
                R.assign(refs.begin() + obj.numLeft, refs.end());
                lb = obj.lb; rb = obj.rb;
            } else if (obj.sah < kBig) {
                const int dim = obj.dim;
                lb = Box(); rb = Box();
                for (const Ref& r : refs) {
                    if (centroid2(r, dim) < obj.pos) { L.push_back(r); lb.grow(r.box); } else { R.push_back(r); rb.grow(r.box); }
                }
            }
            if (L.empty() || R.empty()) {
                // no usable split (all centroids coincide): median cut in storage order
                L.assign(refs.begin(), refs.begin() + n / 2);
                R.assign(refs.begin() + n / 2, refs.end());
                lb = Box(); rb = Box();
                for (const Ref& r : L) lb.grow(r.box);
                for (const Ref& r : R) rb.grow(r.box);
            }
        }
        duplicates += (long long)L.size() + (long long)R.size() - n;
        std::vector<Ref>().swap(refs);      // release before recursing
        tmpBoxes.clear(); tmpBoxes.shrink_to_fit();

        Node* node = ar->newNode();
        node->box = box;
        bool spawned = false;
        if (n >= kParallelThreshold) {
            int cur = activeThreads.load();
            while (cur < maxThreads && !activeThreads.compare_exchange_weak(cur, cur + 1)) {}
            spawned = cur < maxThreads;
        }
        if (spawned) {
            Node* rightNode = nullptr;
            std::thread t([&]() {
                Arena* a2 = newArena();
                rightNode = build(a2, R, rb, depth + 1);
                activeThreads--;
            });
            node->child[0] = build(ar, L, lb, depth + 1);
            t.join();
            node->child[1] = rightNode;
        } else {
            node->child[0] = build(ar, L, lb, depth + 1);
            node->child[1] = build(ar, R, rb, depth + 1);
        }
        return node;
    }
};

}  // namespace

// ---------------------------------------------------------------------------------------------
// flatten to Compact2 + C ABI
// ---------------------------------------------------------------------------------------------
struct b2rt_host_bvh {
    std::vector<float> nodes;       // 16 floats per node
    std::vector<float> woop;        // 4 floats per row
    std::vector<int32_t> triIndex;  // one per Woop row
    int inner = 0, leaves = 0, refs = 0, maxDepth = 0;
    float sah = 0.0f;
};

namespace {

inline float rcpOrZero(float a) { return (a != 0.0f) ? 1.0f / a : 0.0f; }

// 3x3 determinant in the operation order of Math.hh:935-940
inline float det3(const float s[3][3])
{
    return s[0][0] * s[1][1] * s[2][2] - s[0][0] * s[1][2] * s[2][1] +
           s[1][0] * s[2][1] * s[0][2] - s[1][0] * s[2][2] * s[0][1] +
           s[2][0] * s[0][1] * s[1][2] - s[2][0] * s[0][2] * s[1][1];
}

// CudaBVH::woopifyTri (CudaBVH.cc:361-380): rows 2, 0, 1 of the inverse of
// [v0-v2 | v1-v2 | (v0-v2)x(v1-v2) | v2], cofactor expansion as Math.hh:962-983.
void woopify(const float* v0, const float* v1, const float* v2, float out[12])
{
    float m[4][4];      // m[r][c]
    const float e0[3] = { v0[0] - v2[0], v0[1] - v2[1], v0[2] - v2[2] };
    const float e1[3] = { v1[0] - v2[0], v1[1] - v2[1], v1[2] - v2[2] };
    const float nn[3] = { e0[1] * e1[2] - e0[2] * e1[1], e0[2] * e1[0] - e0[0] * e1[2], e0[0] * e1[1] - e0[1] * e1[0] };
    for (int r = 0; r < 3; r++) { m[r][0] = e0[r]; m[r][1] = e1[r]; m[r][2] = nn[r]; m[r][3] = v2[r]; }
    m[3][0] = m[3][1] = m[3][2] = 0.0f; m[3][3] = 1.0f;
    float inv[4][4];
    float d = 0.0f, si = 1.0f;
    for (int i = 0; i < 4; i++) {
        float sj = si;
        for (int j = 0; j < 4; j++) {
            float sub[3][3];
            for (int k = 0; k < 3; k++)
                for (int l = 0; l < 3; l++)
                    sub[k][l] = m[(k < j) ? k : k + 1][(l < i) ? l : l + 1];
            const float dd = det3(sub) * sj;
            inv[i][j] = dd;
            d += dd * m[j][i];
            sj = -sj;
        }
        si = -si;
    }
    const float rd = rcpOrZero(d);
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) inv[i][j] = inv[i][j] * rd * 4.0f;
    out[0] = inv[2][0]; out[1] = inv[2][1]; out[2] = inv[2][2]; out[3] = -inv[2][3];
    for (int c = 0; c < 4; c++) { out[4 + c] = inv[0][c]; out[8 + c] = inv[1][c]; }
    if (out[0] == 0.0f) out[0] = 0.0f;      // -0.0f would read as the leaf terminator (CudaBVH.cc:316-317)
}

void flatten(const Builder& b, const Node* root, b2rt_host_bvh& out)
{
    struct Item { const Node* n; int parent; int which; int depth; float prob; };
    std::vector<Item> stack;
    Node fakeRoot, emptyLeaf;
    if (root->leaf()) {     // the layout needs an inner root (createCompact asserts two children)
        fakeRoot.box = root->box;
        fakeRoot.child[0] = const_cast<Node*>(root);
        fakeRoot.child[1] = &emptyLeaf;
        root = &fakeRoot;
    }
    const float rootArea = root->box.area();
    stack.push_back({ root, -1, 0, 0, 1.0f });
    const uint32_t term = 0x80000000u;
    float termF; memcpy(&termF, &term, 4);
    while (!stack.empty()) {
        const Item it = stack.back(); stack.pop_back();
        int ref;
        out.maxDepth = std::max(out.maxDepth, it.depth);
        if (it.n->leaf()) {
            ref = ~(int)(out.woop.size() / 4);
            for (int i = 0; i < it.n->numTris; i++) {
                const int tri = it.n->tris[i];
                float w[12];
                woopify(b.vtx(tri, 0), b.vtx(tri, 1), b.vtx(tri, 2), w);
                out.woop.insert(out.woop.end(), w, w + 12);
                out.triIndex.push_back(tri); out.triIndex.push_back(0); out.triIndex.push_back(0);
            }
            for (int k = 0; k < 4; k++) out.woop.push_back(termF);
            out.triIndex.push_back(0);
            out.leaves++; out.refs += it.n->numTris;
            out.sah += it.prob * (float)it.n->numTris;
        } else {
            const int idx = (int)(out.nodes.size() / 16);
            ref = idx * 4;
            const Box& b0 = it.n->child[0]->box; const Box& b1 = it.n->child[1]->box;
            const float rec[12] = { b0.lo[0], b0.hi[0], b0.lo[1], b0.hi[1], b1.lo[0], b1.hi[0], b1.lo[1], b1.hi[1],
                                    b0.lo[2], b0.hi[2], b1.lo[2], b1.hi[2] };
            out.nodes.insert(out.nodes.end(), rec, rec + 12);
            out.nodes.insert(out.nodes.end(), 4, 0.0f);
            out.inner++;
            out.sah += it.prob * 2.0f;
            const float pa = it.n->box.area();
            for (int c = 1; c >= 0; c--) {      // child 0 is popped first: it sits right after its parent
                const float cp = (it.prob > 0.0f && pa > 0.0f) ? it.prob * it.n->child[c]->box.area() / pa : 0.0f;
                stack.push_back({ it.n->child[c], idx, c, it.depth + 1, cp });
            }
        }
        if (it.parent >= 0) memcpy(&out.nodes[(size_t)it.parent * 16 + 12 + it.which], &ref, 4);
    }
    (void)rootArea;
}

}  // namespace

extern "C" {

int b2rt_bvh_build(const float* verts, int numVerts, const int32_t* tris, int numTris, const b2rt_build_params* params,
                   b2rt_host_bvh** out)
{
    if (!out) return B2RT_ERR_ARG;
    *out = nullptr;
    if (!verts || !tris || numVerts <= 0 || numTris <= 0) return B2RT_ERR_ARG;
    for (size_t i = 0; i < (size_t)numTris * 3; i++) if (tris[i] < 0 || tris[i] >= numVerts) return B2RT_ERR_ARG;
    Builder b;
    b.verts = verts; b.tris = tris; b.numTris = numTris;
    b.alpha = (params && params->split_alpha > 0.0f) ? params->split_alpha : 1.0e-5f;
    b.minLeaf = (params && params->min_leaf > 0) ? params->min_leaf : 1;
    b.maxLeaf = (params && params->max_leaf > 0) ? params->max_leaf : 8;
    b.spatial = params ? (params->spatial_splits != 0) : true;
    int hw = (int)std::thread::hardware_concurrency();
    if (hw < 1) hw = 1;
    b.maxThreads = (params && params->threads > 0) ? params->threads : hw;

    std::vector<Ref> refs((size_t)numTris);
    Box rootBox;
    for (int i = 0; i < numTris; i++) {
        refs[i].tri = i;
        for (int k = 0; k < 3; k++) refs[i].box.grow(b.vtx(i, k));
        rootBox.grow(refs[i].box);
    }
    b.minOverlap = rootBox.area() * b.alpha;
    Arena* ar = b.newArena();
    Node* root = b.build(ar, refs, rootBox, 0);
    b2rt_host_bvh* h = new b2rt_host_bvh();
    h->nodes.reserve((size_t)numTris * 16);
    h->woop.reserve((size_t)numTris * 16);
    flatten(b, root, *h);
    *out = h;
    return B2RT_OK;
}

void b2rt_bvh_free(b2rt_host_bvh* bvh) { delete bvh; }

int b2rt_bvh_arrays(const b2rt_host_bvh* bvh, const void** nodes, size_t* nodeBytes, const void** woop, size_t* woopBytes,
                    const int32_t** triIndex, size_t* triIndexBytes)
{
    if (!bvh) return B2RT_ERR_ARG;
    if (nodes) *nodes = bvh->nodes.data();
    if (nodeBytes) *nodeBytes = bvh->nodes.size() * 4;
    if (woop) *woop = bvh->woop.data();
    if (woopBytes) *woopBytes = bvh->woop.size() * 4;
    if (triIndex) *triIndex = bvh->triIndex.data();
    if (triIndexBytes) *triIndexBytes = bvh->triIndex.size() * 4;
    return B2RT_OK;
}

int b2rt_bvh_stats(const b2rt_host_bvh* bvh, int* inner, int* leaves, int* refs, float* sah, int* maxDepth)
{
    if (!bvh) return B2RT_ERR_ARG;
    if (inner) *inner = bvh->inner;
    if (leaves) *leaves = bvh->leaves;
    if (refs) *refs = bvh->refs;
    if (sah) *sah = bvh->sah;
    if (maxDepth) *maxDepth = bvh->maxDepth;
    return B2RT_OK;
}

// bvhcache format: S32 layout, then nodes / triWoop / triIndex each as S64 byte size + bytes, little endian
int b2rt_bvh_save(const b2rt_host_bvh* bvh, const char* path)
{
    if (!bvh || !path) return B2RT_ERR_ARG;
    FILE* f = fopen(path, "wb");
    if (!f) return B2RT_ERR_ARG;
    const int32_t layout = B2RT_LAYOUT_COMPACT2;
    bool ok = fwrite(&layout, 4, 1, f) == 1;
    const void* ptr[3] = { bvh->nodes.data(), bvh->woop.data(), bvh->triIndex.data() };
    const int64_t size[3] = { (int64_t)bvh->nodes.size() * 4, (int64_t)bvh->woop.size() * 4, (int64_t)bvh->triIndex.size() * 4 };
    for (int i = 0; i < 3 && ok; i++) ok = fwrite(&size[i], 8, 1, f) == 1 && (size[i] == 0 || fwrite(ptr[i], 1, (size_t)size[i], f) == (size_t)size[i]);
    fclose(f);
    return ok ? B2RT_OK : B2RT_ERR_ARG;
}

int b2rt_bvh_load(const char* path, b2rt_host_bvh** out)
{
    if (!path || !out) return B2RT_ERR_ARG;
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return B2RT_ERR_ARG;
    std::unique_ptr<b2rt_host_bvh> h(new b2rt_host_bvh());
    int32_t layout = -1;
    bool ok = fread(&layout, 4, 1, f) == 1;
    int rc = B2RT_OK;
    if (ok && layout != B2RT_LAYOUT_COMPACT2) { ok = false; rc = B2RT_ERR_LAYOUT; }
    int64_t size = 0;
    if (ok && (ok = fread(&size, 8, 1, f) == 1 && size >= 0 && size % 64 == 0)) { h->nodes.resize((size_t)size / 4); ok = size == 0 || fread(h->nodes.data(), 1, (size_t)size, f) == (size_t)size; }
    if (ok && (ok = fread(&size, 8, 1, f) == 1 && size >= 0 && size % 16 == 0)) { h->woop.resize((size_t)size / 4); ok = size == 0 || fread(h->woop.data(), 1, (size_t)size, f) == (size_t)size; }
    if (ok && (ok = fread(&size, 8, 1, f) == 1 && size >= 0 && size % 4 == 0)) { h->triIndex.resize((size_t)size / 4); ok = size == 0 || fread(h->triIndex.data(), 1, (size_t)size, f) == (size_t)size; }
    fclose(f);
    if (!ok) return rc ? rc : B2RT_ERR_ARG;
    h->inner = (int)(h->nodes.size() / 16);
    *out = h.release();
    return B2RT_OK;
}

int b2rt_tri_normals(const float* verts, const int32_t* tris, int numTris, float* normals)
{
    if (!verts || !tris || !normals || numTris < 0) return B2RT_ERR_ARG;
    for (int i = 0; i < numTris; i++) {
        const float* a = verts + 3 * (size_t)tris[3 * (size_t)i];
        const float* b = verts + 3 * (size_t)tris[3 * (size_t)i + 1];
        const float* c = verts + 3 * (size_t)tris[3 * (size_t)i + 2];
        const float e1[3] = { b[0] - a[0], b[1] - a[1], b[2] - a[2] }, e2[3] = { c[0] - a[0], c[1] - a[1], c[2] - a[2] };
        const float n[3] = { e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0] };
        float l = 0.0f;
        l += n[0] * n[0]; l += n[1] * n[1]; l += n[2] * n[2];
        const float s = 1.0f * rcpOrZero(sqrtf(l));
        normals[3 * (size_t)i] = n[0] * s; normals[3 * (size_t)i + 1] = n[1] * s; normals[3 * (size_t)i + 2] = n[2] * s;
    }
    return B2RT_OK;
}

}  // extern "C"
