/*
 * b2rt_mg.h -- multi-GPU entry points of libb2rt.so: the BVH replicated on every GPU of one box, ray batches sharded.
 *
 * The reference is single-GPU (no NCCL, no peer access: `grep -i "nccl\|mpi\|peer"` over its src/ finds nothing), so these
 * have no reference interface to cite; the contract is BASELINE.json's north_star ("replicating the BVH on every GPU and
 * sharding ray batches ... NCCL over NVLink is used only to broadcast the scene and to gather hit buffers and stats") and
 * SURVEY.md 8(b)/(e).  Rays are independent, so there is NO collective inside a trace:
 *   b2rt_mg_create          one b2rt_ctx + one stream per device, one NCCL communicator over them (ncclCommInitAll; one
 *                           host thread drives all devices, NCCL calls are grouped)
 *   b2rt_mg_broadcast_bvh   the BVH goes to the first device ONCE; its private forms (re-laid nodes, padded Woop rows, and --
 *                           when the default policy would use it -- the 8-wide compressed form) are built there ONCE and
 *                           ncclBroadcast straight into the other contexts' private arrays: no host round trip per rank
 *   b2rt_mg_trace           each device traces its shard (device pointers on that device); optionally the 16-byte hit
 *                           records of all shards are gathered into every device's `gathered` buffer (one ncclBroadcast per
 *                           shard inside one group = an all-gather of unequal parts) and the hit counts summed with
 *                           ncclAllReduce; times are CUDA events per device, the reported trace time is the max over devices
 *   b2rt_mg_trace_host      the same for a caller holding ONE host ray buffer: contiguous slices of ceil(n / g) rays go up,
 *                           are traced and come back into the matching slice of the host result buffer
 * NCCL is bound at run time (dlopen of libnccl.so.2): libb2rt.so has no link-time dependency on it, and a process that
 * already carries NCCL (torch's) shares that copy.  Everything returns 0 or a negative b2rt_status; the message of the last
 * failure is b2rt_mg_last_error(mg) (mg == NULL: the last failure of b2rt_mg_create).
 */
#ifndef B2RT_MG_H
#define B2RT_MG_H

#include "b2rt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2rt_mg b2rt_mg;

typedef struct {
    int64_t rays;               /* launched rays, all devices */
    int64_t hits;               /* records with id >= 0, all devices (ncclAllReduce sum of the per-device counts) */
    float   ms_trace_max;       /* max over devices of the trace launch's CUDA-event time: the figure Mrays/s is quoted on */
    float   ms_trace_min;
    float   ms_with_gather_max; /* max over devices, first trace event to the end of the gather + all-reduce (0 without gather) */
    float   ms_per_device[16];  /* trace time of each device */
} b2rt_mg_stats;

/* devices == NULL: devices 0 .. num_devices-1.  At most 16 devices. */
int  b2rt_mg_create(int num_devices, const int* devices, b2rt_mg** out);
void b2rt_mg_destroy(b2rt_mg* mg);
const char* b2rt_mg_last_error(const b2rt_mg* mg);
int  b2rt_mg_num_devices(const b2rt_mg* mg);
/* the per-device context (for b2rt_raygen_*, b2rt_morton_sort ... on that device) and its stream (a cudaStream_t) */
b2rt_ctx* b2rt_mg_context(b2rt_mg* mg, int rank);
void* b2rt_mg_stream(b2rt_mg* mg, int rank);

/* arrays in host memory, or (src_is_device) on the FIRST device of the group.  Synchronous. */
int b2rt_mg_broadcast_bvh(b2rt_mg* mg, const void* nodes, size_t node_bytes, const void* woop, size_t woop_bytes,
                          const int32_t* tri_index, size_t tri_index_bytes, int layout, int src_is_device);

/* rays_dev[r], results_dev[r]: device r's shard (num_rays[r] rays, may be 0).  gathered_dev: NULL, or per device a buffer of
 * sum(num_rays) 16-byte records that receives all shards in rank order.  Synchronous (it returns times). */
int b2rt_mg_trace(b2rt_mg* mg, const void* const* rays_dev, void* const* results_dev, const int64_t* num_rays,
                  unsigned flags, void* const* gathered_dev, b2rt_mg_stats* stats);

/* one host ray buffer in, one host result buffer out; slice r = [r * ceil(n/g), min(n, (r+1) * ceil(n/g))) */
int b2rt_mg_trace_host(b2rt_mg* mg, const void* rays_host, void* results_host, int64_t num_rays, unsigned flags,
                       b2rt_mg_stats* stats);

#ifdef __cplusplus
}
#endif
#endif /* B2RT_MG_H */
