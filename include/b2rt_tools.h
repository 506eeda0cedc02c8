/*
 * b2rt_tools.h -- instrumentation and inspection entry points of libb2rt.so that are NOT part of the drop-in
 * boundary (include/b2rt.h): SIMD-efficiency counters and the launch timeline of the instrumented kernel, the
 * canonical work counters behind bench.py's roofline, and the host-side build of the private 8-wide form so that
 * it can be inspected without a GPU.  Used by bench.py, tools/ and tests/ only.
 */
#ifndef B2RT_TOOLS_H
#define B2RT_TOOLS_H

#include "b2rt.h"

#ifdef __cplusplus
extern "C" {
#endif

/* SIMD-efficiency counters of the instrumented kernel (README.md:46-81 TRV / IST columns:
 * mean active lanes / 32 over node-loop and triangle-loop iterations). */
typedef struct {
    uint64_t node_iters;        /* warp-level node-loop iterations */
    uint64_t node_lanes;        /* active lanes summed over them */
    uint64_t tri_iters;         /* warp-level triangle-loop iterations */
    uint64_t tri_lanes;
    uint64_t fetches;           /* warp-level ray fetches (atomics on the work counter) */
    uint64_t rays;
    /* launch timeline from the global timer (binary kernel): nanoseconds after the first warp started */
    uint64_t warps;             /* persistent warps that ran */
    uint64_t ns_pool_empty;     /* first warp found the work counter past the last ray */
    uint64_t ns_mean_exit;      /* mean warp exit time */
    uint64_t ns_last_exit;      /* last warp exit = length of the launch; (last - mean) / last = idle tail */
    uint64_t ns_exit_p50, ns_exit_p90, ns_exit_p99;    /* percentiles of the per-warp exit times */
    /* the warp that left last: when it fetched its last rays, and the loop iterations it ran from then on --
     * the launch cannot end before that one bundle's dependent chain of node and triangle steps has */
    uint64_t ns_straggler_fetch, straggler_node_iters, straggler_tri_iters;
    uint64_t shares;            /* any-hit kernels: stack entries handed to idle lanes after the pool ran dry */
} b2rt_trace_stats;

/* counters of the last B2RT_TRACE_STATS launch (synchronises) */
int b2rt_get_trace_stats(b2rt_ctx* ctx, b2rt_trace_stats* out);

/* Algorithmic work of a batch: per-ray counters of the CANONICAL traversal (near-first ordered
 * stack traversal of the Compact2 arrays, Woop test, closest-hit pruning or any-hit exit, no
 * speculation, no postponement, IEEE division) summed over the batch -- the N_n, N_t, N_l, N_u,
 * N_v, h of SURVEY.md 8(d) from which bench.py computes the roofline's algorithmic bytes.  A plain
 * one-thread-per-ray kernel, not the production tracer.  Synchronous. */
typedef struct {
    uint64_t rays, nodes, tris, leaves, ustage, vstage, hits, max_stack;
    uint64_t max_chain;         /* largest (node visits + triangle tests) of a single ray: the serial tail of a launch */
} b2rt_work_counters;
int b2rt_count_work(b2rt_ctx* ctx, const void* rays_dev, int64_t num_rays, int any_hit,
                    b2rt_work_counters* out, void* stream);

/* Private 8-wide compressed form of a Compact2 BVH (csrc/wide_bvh.h): what b2rt_set_bvh builds for the wide
 * traversal kernel (b2rt_trace_config.variant >= 100), exposed so it can be built and inspected without a GPU.
 * 80-byte nodes, re-packed Woop rows (3 per triangle, +2 padding rows), one original triangle id per triangle.
 * stats = { nodes, triangles, depth }. */
typedef struct b2rt_wide_bvh b2rt_wide_bvh;
int  b2rt_wide_build(const void* nodes, size_t node_bytes, const void* woop, size_t woop_bytes,
                     const int32_t* tri_index, size_t tri_index_bytes, b2rt_wide_bvh** out);
void b2rt_wide_free(b2rt_wide_bvh* w);
int  b2rt_wide_arrays(const b2rt_wide_bvh* w, const void** nodes, size_t* node_bytes, const void** woop,
                      size_t* woop_bytes, const int32_t** tri_index, size_t* tri_index_bytes, int stats[3]);

#ifdef __cplusplus
}
#endif
#endif /* B2RT_TOOLS_H */
