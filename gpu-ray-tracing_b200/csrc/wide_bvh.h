// wide_bvh.h -- private 8-wide compressed re-layout of a caller's Compact2 BVH.
//
// b2rt_set_bvh keeps PRIVATE copies of the caller's arrays (include/b2rt.h).  Besides the re-ordered
// 64-byte child-pair nodes (api.cu:relayoutNodes) the context can hold a second private form built
// from the same Compact2 arrays: the binary tree collapsed to 8-wide nodes with 8-bit quantised
// child boxes, 80 bytes per node (the layout of Ylitie, Karras, Laine, "Efficient Incoherent Ray
// Traversal on GPUs Through Compressed Wide BVHs", HPG 2017), and the Woop rows re-packed so that
// the triangles under one wide node are contiguous (three rows each, no terminator rows).  The
// reference sketches wider layouts only as an enum (kernels/CudaTracerKernels.hh:117-127); this is
// SURVEY.md 8(f) row 4, done inside the library so the caller's contract -- Compact2 in, RayResult
// out, ids through the caller's triIndex -- does not change.  Triangle tests run on the caller's own
// Woop rows, so t, u, v are bit-identical to the binary path; only the ORDER in which equal-t
// candidates are met can differ.
//
// Node (5 x 16 bytes):
//   q0: float p.x, p.y, p.z ; bytes { ex, ey, ez, imask }   ex = biased exponent of the grid step, +15
//   q1: uint childBase ; uint triBase ; bytes meta[0..3] ; bytes meta[4..7]
//   q2: qlo_x[0..7], qlo_y[0..7]      q3: qlo_z[0..7], qhi_x[0..7]      q4: qhi_y[0..7], qhi_z[0..7]
// meta[i]: 0 = empty slot; inner child: 0b001_11sss (sss = slot i), also bit i of imask; leaf child:
// high 3 bits = triangle count in unary (1, 3, 7), low 5 bits = first triangle relative to triBase.
// Child box i = p + 2^e * q, e = byte - 142 (see trace_wide.cu for the decode).  The boxes are conservative supersets of the
// binary tree's WITH the kernel's rounding: each plane is quantised with 1/64 of a cell of slack (host/wide_bvh.cpp).
// Which binary nodes are folded into one wide node is decided by a dynamic programme over the binary tree that minimises
// the SAH cost of the result (Ylitie et al., section 3.1); subtrees with at most three triangle references become one leaf slot.
#pragma once

#include <stddef.h>
#include <stdint.h>

#include <vector>

namespace b2rt {

struct Wide8 {
    std::vector<uint32_t> nodes;       // 20 words per node
    std::vector<float> woop;           // 12 floats per triangle (3 rows), triangles of a node contiguous
    std::vector<int32_t> triIndex;     // original triangle id per private triangle
    // statistics
    int numNodes = 0, numTris = 0, maxDepth = 0;
    double avgChildren = 0.0;
};

// Collapse.  nodes2: Compact2 nodes (16 floats each, caller's order, root = node 0); woop / triIndex: the
// caller's rows.  Returns false (with a message) on malformed input.
bool buildWide8(const float* nodes2, size_t numNodes2, const float* woop, const int32_t* triIndex, size_t numRows,
                Wide8& out, const char** error, bool greedy = false);

// Structural check used by the CPU tests: every triangle reference of the binary tree appears exactly once, every
// wide node is reachable exactly once, indices are in range.  (That every quantised child box contains the binary
// tree's box with at least half a margin to spare is checked by buildWide8 itself while it emits a node -- it is the
// only place that still knows which binary box a slot came from -- and makes the build fail.)
bool validateWide8(const float* nodes2, size_t numNodes2, const float* woop, const int32_t* triIndex, size_t numRows,
                   const Wide8& w, const char** error);

}  // namespace b2rt
