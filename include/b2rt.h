/*
 * b2rt.h -- C ABI of the Blackwell-native ray-casting core (libb2rt.so, sm_100a).
 *
 * Drop-in boundary for ONE hot path of yanrun000/gpu-ray-tracing: the batched closest-hit /
 * any-hit trace of a RayBuffer against a CudaBVH, with the ray-generation, ray-sort and
 * hit-count kernels either side of it.  Plain pointers and sizes only; every entry point
 * returns 0 on success and a negative b2rt_status otherwise, never exits the process (the
 * reference's convention is fprintf + exit(-1): src/rt/cutil_inline_runtime.h:33-42,
 * src/framework/base/Defs.cc:41-54) and leaves a message in b2rt_last_error().
 *
 * Each declaration cites the reference interface it replaces (paths under the reference's
 * src/).  Record layouts are the reference's, byte for byte:
 *   Ray        32 B  {float3 origin, float tmin, float3 direction, float tmax}   rt/Util.hh:64-73
 *   RayResult  16 B  {int32 id, float t, int32 padA, int32 padB}                  rt/Util.hh:79-89
 *              padA/padB are never written by the reference (kernels/CudaTracerKernels.hh:197);
 *              with B2RT_TRACE_WRITE_BARY this library stores barycentric u,v there as floats.
 *   node       64 B  child-pair node of BVHLayout_Compact2                        rt/cuda/CudaBVH.hh:41-44
 *   Woop tri   48 B  three float4 rows, leaf terminated by a row whose .x = 0x80000000   CudaBVH.hh:47-50
 *   triIndex    4 B  per Woop row: original triangle id                           CudaBVH.cc:319-327
 *
 * Ownership: the caller owns every device array it passes (as RayBuffer and CudaBVH do in the
 * reference, RayBuffer.cc:42-81, CudaBVH.cc:101-109).  b2rt_set_bvh keeps a PRIVATE re-laid
 * copy of the node array (top levels first, for shared-memory staging) and plain copies of the
 * Woop / triIndex arrays, so the caller's BVH arrays may be freed after it returns.
 * Threading: one b2rt_ctx per device; a ctx is not thread-safe, different ctxs are independent.
 * All *_dev pointers are device pointers on the ctx's device; `stream` is a cudaStream_t cast to
 * void* (NULL = default stream).  Calls are asynchronous on `stream` unless stated.
 */
#ifndef B2RT_H
#define B2RT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2rt_ctx b2rt_ctx;

typedef enum {
    B2RT_OK             =  0,
    B2RT_ERR_ARG        = -1,   /* bad argument */
    B2RT_ERR_CUDA       = -2,   /* CUDA runtime error (message has the cudaError string) */
    B2RT_ERR_NO_BVH     = -3,   /* "CudaTracer: No BVH!"               rt/cuda/CudaTracer.cc:128 */
    B2RT_ERR_LAYOUT     = -4,   /* "CudaTracer: Incorrect BVH layout!" rt/cuda/CudaTracer.cc:130 */
    B2RT_ERR_NO_DEVICE  = -5,   /* no usable sm_100 device */
    B2RT_ERR_ALLOC      = -6
} b2rt_status;

/* rt/kernels/CudaTracerKernels.hh:117-127 -- only Compact2 is accepted (CudaTracer.cc:103) */
#define B2RT_LAYOUT_COMPACT2 5
/* rt/kernels/CudaTracerKernels.hh:110 */
#define B2RT_ENTRYPOINT_SENTINEL 0x76543210

/* b2rt_trace flags */
#define B2RT_TRACE_ANY_HIT     1u  /* !RayBuffer::getNeedClosestHit()              CudaTracer.cc:172 */
#define B2RT_TRACE_WRITE_BARY  2u  /* store u,v in RayResult.padA/padB             (superset) */
#define B2RT_TRACE_WATERTIGHT  8u  /* edge-tolerant Woop test + order-independent closest hit: no leaks through shared edges, the same
                                      record for a ray wherever it sits in the batch.  Off by default: the reference's test is exact
                                      (kepler_dynamic_fetch.cu:349-364) and so is parity with it.  Implies WRITE_BARY. */
#define B2RT_TRACE_STATS       4u  /* instrumented kernel: fills b2rt_trace_stats (include/b2rt_tools.h; slower) */

/* Tunables of the traversal kernel; zero-initialise and override what you need. */
typedef struct {
    int top_nodes;              /* nodes staged in shared memory (0 = none, -1 = default) */
    int fetch_threshold;        /* refill dead lanes when fewer than this are alive
                                   (kepler_dynamic_fetch.cu:48 uses 20); -1 = default: 20, 24 in the 8-wide kernel */
    int blocks_per_sm;          /* persistent CTAs per SM; -1 = from occupancy */
    int variant;                /* kernel variant id; -1 = chosen from the batch size and the size of the BVH (trace.cu:launchTrace) */
    int l2_prefetch;            /* reserved (was: pull the BVH into L2 at launch start; measured no gain, removed); ignored */
    unsigned tune;              /* B2RT_TUNE_* bits that switch single design decisions OFF for A/B measurements; 0 = as shipped */
} b2rt_trace_config;
#define B2RT_TUNE_NO_STREAM      1u  /* rays / records through the normal cache policy instead of evict-first (ld.cs / st.cs);
                                         honoured by -DB2RT_EXPLORE builds only: shipped kernels always stream */
#define B2RT_TUNE_GREEDY_WIDE    4u  /* 8-wide form: round-1 greedy top-down collapse instead of the SAH-optimal one; takes effect
                                        at the next b2rt_set_bvh */
#define B2RT_TUNE_DEFAULT_CARVEOUT 16u /* leave the shared-memory / L1 split to the driver instead of asking for the resident CTAs' need;
                                        takes effect for kernels not launched yet on this context */
#define B2RT_TUNE_L2_WINDOW      8u  /* (opt-in) 8-wide kernel: persisting-L2 access window over the private wide arrays when they fit
                                        the device's set-aside; sets the device-wide cudaLimitPersistingL2CacheSize */

/* ---- context ---------------------------------------------------------------------------- */
/* replaces CudaModule::staticInit + CudaTracer::CudaTracer (CudaTracer.cc:67-72) */
int  b2rt_create(int device, b2rt_ctx** out);
void b2rt_destroy(b2rt_ctx* ctx);
/* message of the last failure on this ctx (ctx == NULL: last failure of b2rt_create) */
const char* b2rt_last_error(const b2rt_ctx* ctx);
/* number of kernels this ctx has launched so far (bench.py's gpu_launches) */
int64_t b2rt_launch_count(const b2rt_ctx* ctx);
int  b2rt_set_trace_config(b2rt_ctx* ctx, const b2rt_trace_config* cfg);
int  b2rt_device_info(const b2rt_ctx* ctx, int* sm_count, int* cc_major, int* cc_minor);

/* ---- BVH -------------------------------------------------------------------------------- */
/* replaces CudaTracer::setBVH + bind_CudaBVHTexture (CudaTracer.hh:49, CudaTracer.cc:142-146,
 * kernels/kepler_dynamic_fetch.cu:417-424).  Synchronous.  layout must be Compact2. */
int b2rt_set_bvh(b2rt_ctx* ctx, const void* nodes, size_t node_bytes, const void* woop,
                 size_t woop_bytes, const int32_t* tri_index, size_t tri_index_bytes, int layout,
                 int src_is_device);
/* the arrays the kernels read (device pointers; the node array is the re-laid private copy) */
int b2rt_get_bvh(const b2rt_ctx* ctx, const void** nodes_dev, size_t* node_bytes,
                 const void** woop_dev, size_t* woop_bytes, const int32_t** tri_index_dev,
                 size_t* tri_index_bytes);

/* ---- trace ------------------------------------------------------------------------------ */
/* replaces CudaTracer::traceBatch + launch_tracingKernel (CudaTracer.cc:119-178,
 * kepler_dynamic_fetch.cu:432-474).  If ms != NULL the launch is bracketed by CUDA events on
 * `stream`, the call synchronises and *ms receives the kernel time in milliseconds (what the
 * reference returns); if ms == NULL the call is asynchronous.  num_rays == 0 returns 0 ms.
 * Asynchronous traces issued on DIFFERENT streams may overlap (the reference allows one trace in flight per device:
 * one g_warpCounter, kepler_dynamic_fetch.cu:58): up to 8 may be in flight per context, so the idle tail of one batch
 * can be filled by the next independent one.  B2RT_TRACE_STATS launches are not safe to overlap. */
int b2rt_trace(b2rt_ctx* ctx, const void* rays_dev, void* results_dev, int64_t num_rays,
               unsigned flags, void* stream, float* ms);
/* end-to-end variant for callers holding HOST buffers (pinned or pageable): H2D copy of the
 * rays, trace, D2H copy of the results, all inside the call.  ms_kernel / ms_total optional. */
int b2rt_trace_host(b2rt_ctx* ctx, const void* rays_host, void* results_host, int64_t num_rays,
                    unsigned flags, float* ms_kernel, float* ms_total);
/* The same in two halves, for callers with several independent batches (a frame's batches, CudaTracer.cc:119-178 called
 * in a loop by Renderer.cc:262-330): _begin enqueues upload, trace and download and returns; up to two batches may be
 * in flight (a third _begin first waits for the oldest), so the second batch's upload overlaps the first one's last
 * kernel and download.  _end waits for everything begun and returns the summed kernel time and the device time from
 * the first upload to the last download.  The host buffers must stay valid, and the results unread, until _end. */
int b2rt_trace_host_begin(b2rt_ctx* ctx, const void* rays_host, void* results_host, int64_t num_rays, unsigned flags);
int b2rt_trace_host_end(b2rt_ctx* ctx, float* ms_kernel, float* ms_total);

/* ---- ray generation --------------------------------------------------------------------- */
/* replaces launch_rayGenPrimaryKernel (ray/RayGenKernels.hh:40-51,107-109; RayGenKernels.cu:79-113).
 * nscreen_to_world is column-major like FW::Mat4f.  Writes w*h rays and both id maps. */
int b2rt_raygen_primary(b2rt_ctx* ctx, const float origin[3], const float nscreen_to_world[16],
                        int w, int h, float max_dist, const int32_t* index_to_pixel_dev,
                        void* rays_dev, int32_t* id_to_slot_dev, int32_t* slot_to_id_dev,
                        void* stream);
/* replaces launch_rayGenAOKernel (RayGenKernels.hh:55-68,110; RayGenKernels.cu:117-227).
 * AO and Diffuse share it (Renderer.cc:268-277).  Writes num_input_rays*num_samples rays. */
int b2rt_raygen_ao(b2rt_ctx* ctx, int first_input_slot, int num_input_rays, int num_samples,
                   float max_dist, uint32_t random_seed, const void* in_rays_dev,
                   const void* in_results_dev, const float* normals_dev, void* out_rays_dev,
                   int32_t* out_id_to_slot_dev, int32_t* out_slot_to_id_dev, void* stream);

/* ---- Morton ray sort -------------------------------------------------------------------- */
/* replaces RayBuffer::mortonSort (ray/RayBuffer.cc:256-324) and its three launchers
 * (ray/RayBufferKernels.hh:112-117) plus the host quicksort (framework/base/Sort.cc:131-165):
 * AABB of origins and end points, 192-bit 6-D Morton key, device radix sort, gather of rays and
 * rebuild of both id maps -- all on the device, in place.  The call synchronises `stream` (it
 * reads one flag back to decide between its two sort paths).  Order inside runs of equal keys is
 * unspecified in the reference (unstable quicksort); here it is slot order (stable). */
int b2rt_morton_sort(b2rt_ctx* ctx, void* rays_dev, int32_t* id_to_slot_dev,
                     int32_t* slot_to_id_dev, int64_t num_rays, float aabb_out[6], void* stream);
/* Sort path: 0 (default) = radix sort by the key's top 56 significant bits, then ties settled by the full key; falls back
 * by itself to 1 = radix sort over all 192 bits when a run of equal prefixes is too long.  Both give the same order.
 * A mode in 64..139 is path 0 with that prefix shift instead of the default 91 (tuning: (147 - shift) / 8 passes).
 * b2rt_sort_fell_back: 1 if the last b2rt_morton_sort went the long way. */
int b2rt_set_sort_mode(b2rt_ctx* ctx, int mode);
int b2rt_sort_fell_back(const b2rt_ctx* ctx);
/* the pieces, for parity tests: AABB (RayBufferKernels.cu:74-140) and keys (:144-179) */
int b2rt_find_aabb(b2rt_ctx* ctx, const void* rays_dev, int64_t num_rays, float aabb_out[6],
                   void* stream);
int b2rt_morton_keys(b2rt_ctx* ctx, const void* rays_dev, int64_t num_rays, const float aabb[6],
                     void* keys_dev /* 28 B each: {int32 oldSlot; uint32 hash[6]} */, void* stream);

/* replaces launch_reorderRaysKernel (RayBufferKernels.hh:81-91,116; RayBufferKernels.cu:183-200): gather rays by
 * the oldSlot of sorted 28-byte MortonKey records and rebuild both id maps.  Out of place. */
int b2rt_reorder_rays(b2rt_ctx* ctx, const void* keys_dev, const void* in_rays_dev,
                      const int32_t* in_slot_to_id_dev, void* out_rays_dev, int32_t* out_id_to_slot_dev,
                      int32_t* out_slot_to_id_dev, int64_t num_rays, void* stream);

/* ---- hit count / reconstruction --------------------------------------------------------- */
/* replaces launch_countHitsKernel (cuda/RendererKernels.hh:77-92, RendererKernels.cu:112-162):
 * number of results with id >= 0.  Synchronous. */
int b2rt_count_hits(b2rt_ctx* ctx, const void* results_dev, int64_t num_rays, int64_t* hits,
                    void* stream);
/* NO counterpart in the reference: added for callers that cannot allocate (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The hit count of b2rt_count_hits
 * without a context and without any allocation: scratch lives in a file-scope device array of the library (33 KB, held by
 * every GPU that loads the library until the process exits), the total is written to the DEVICE word hits_dev, and the
 * call only enqueues an 8-byte memset and one kernel on `stream` (current device; asynchronous, no synchronisation, no
 * cudaMalloc / cudaFree).  num_rays above B2RT_COUNT_HITS_STATIC_MAX is refused with B2RT_ERR_ARG; errors of this
 * context-less call are read with b2rt_last_error(NULL).  All calls on a device share the scratch: calls that overlap on
 * the GPU (different streams) race; calls on one stream are safe.  csrc/static_scratch.cu. */
#define B2RT_COUNT_HITS_STATIC_MAX (16 * 1024 * 1024)
int b2rt_count_hits_static(const void* results_dev, int64_t num_rays, int64_t* hits_dev, void* stream);
/* replaces launch_reconstructKernel (RendererKernels.hh:54-70, RendererKernels.cu:60-108):
 * ABGR pixel from the batch's hit ids.  mode: 0 primary, 1 AO, 2 diffuse. */
int b2rt_reconstruct(b2rt_ctx* ctx, int num_rays_per_primary, int first_primary, int num_primary,
                     int mode, const int32_t* primary_slot_to_id_dev, const void* primary_results_dev,
                     const int32_t* batch_id_to_slot_dev, const void* batch_results_dev,
                     const uint32_t* tri_material_color_dev, const uint32_t* tri_shaded_color_dev,
                     uint32_t* pixels_dev, void* stream);

/* ---- host-side helpers (no GPU needed) --------------------------------------------------- */
/* replaces PixelTable::recalculate (ray/PixelTable.cc:70-161); pixel_to_index may be NULL */
int b2rt_pixel_table(int w, int h, int32_t* index_to_pixel, int32_t* pixel_to_index);
/* replaces CameraControls::getWorldToClip + Renderer::beginFrame's matrix (Renderer.cc:126-129) */
int b2rt_camera_matrices(const float pos[3], const float forward[3], const float up[3], float fov,
                         float znear, float zfar, int w, int h, float origin[3],
                         float nscreen_to_world[16]);

/* Host BVH: SBVH build (rt/bvh/SplitBVHBuilder.cc) + flatten to Compact2 with Woop transform
 * (rt/cuda/CudaBVH.cc:270-380).  Same algorithm family and the same output layout as the
 * reference; the tree itself is not bit-identical to the reference builder's. */
typedef struct b2rt_host_bvh b2rt_host_bvh;
typedef struct {
    float split_alpha;          /* BVH::BuildParams::splitAlpha, default 1e-5 (bvh/BVH.hh:79) */
    int   min_leaf, max_leaf;   /* Platform leaf preferences, default 1 / 8 (Renderer.cc:54) */
    int   threads;              /* 0 = hardware concurrency */
    int   spatial_splits;       /* 1 = SBVH, 0 = object splits only */
} b2rt_build_params;
int  b2rt_bvh_build(const float* verts, int num_verts, const int32_t* tris, int num_tris,
                    const b2rt_build_params* params, b2rt_host_bvh** out);
void b2rt_bvh_free(b2rt_host_bvh* bvh);
int  b2rt_bvh_arrays(const b2rt_host_bvh* bvh, const void** nodes, size_t* node_bytes,
                     const void** woop, size_t* woop_bytes, const int32_t** tri_index,
                     size_t* tri_index_bytes);
int  b2rt_bvh_stats(const b2rt_host_bvh* bvh, int* inner_nodes, int* leaves, int* tri_refs,
                    float* sah_cost, int* max_depth);
/* bvhcache file format (Renderer.cc:176-211; CudaBVH.cc:79-97,113-116; Buffer.cc:327-361) */
int  b2rt_bvh_save(const b2rt_host_bvh* bvh, const char* path);
int  b2rt_bvh_load(const char* path, b2rt_host_bvh** out);
/* The name Renderer::getCudaBVH gives the cache file of a BVH: <dir>/%08x.dat with
 * key = hashBits(scene.hash(), platform.computeHash(), buildParams.computeHash(), layout)  (Renderer.cc:176-186; Scene.cc:94-102,
 * bvh/Platform.hh:69, bvh/BVH.hh:82-85, framework/base/Hash.hh:181-196, Hash.cc:34-105), so that a bvhcache directory can be
 * shared with a reference build.  verts / tris: the arrays of the reference's Scene for the mesh (vertices in the importer's
 * order); tri_diffuse_rgba: 4 floats per triangle, or NULL for the default material (0.75 grey, 3d/Mesh.hh:92). */
int  b2rt_bvhcache_key(const float* verts, int num_verts, const int32_t* tris, int num_tris, const float* tri_diffuse_rgba,
                       float split_alpha, int min_leaf, int max_leaf, int layout, uint32_t* key);
int  b2rt_bvhcache_path(const char* dir, uint32_t key, char* out, size_t cap);
/* per-triangle normals as Scene does (rt/Scene.cc:72) */
int  b2rt_tri_normals(const float* verts, const int32_t* tris, int num_tris, float* normals);
/* The Woop rows of a Compact2 BVH computed on the device: replaces the per-triangle part of CudaBVH::createCompact /
 * woopifyTri (CudaBVH.cc:270-380; matrix inverse of Math.hh:935-983), bit-identical to the host computation.
 * All pointers are device memory.  verts: xyz per vertex; vtx_index: three vertex indices per triangle; reference r
 * (r < num_refs) is triangle ref_tri[r] and occupies rows ref_row[r] .. ref_row[r]+2 of `woop` (num_rows 16-byte rows);
 * every row no reference covers becomes a leaf terminator (x = 0x80000000, CudaBVH.hh:47-50). */
int  b2rt_woopify(b2rt_ctx* ctx, const float* verts, const int32_t* vtx_index, const int32_t* ref_row, const int32_t* ref_tri,
                  int64_t num_refs, void* woop, int64_t num_rows, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2RT_H */
