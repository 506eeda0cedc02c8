// common.cuh -- shared declarations of libb2rt (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <map>
#include <string>
#include <utility>

#include "../../include/b2rt.h"
#include "../../include/b2rt_tools.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libb2rt is written for sm_100a (B200) only"
#endif

namespace b2rt {

// rt/Util.hh:64-89 -- record layouts (ABI)
struct Ray       { float ox, oy, oz, tmin, dx, dy, dz, tmax; };
struct RayResult { int32_t id; float t; int32_t padA, padB; };
struct MortonKey { int32_t oldSlot; uint32_t hash[6]; };       // ray/RayBufferKernels.hh:46-50
static_assert(sizeof(Ray) == 32 && sizeof(RayResult) == 16 && sizeof(MortonKey) == 28, "ABI");

constexpr int kSentinel = B2RT_ENTRYPOINT_SENTINEL;
constexpr int kTimelineWarps = 8192;    // per-warp exit times kept by the instrumented kernel (statsDev[16 ...])
constexpr int kWideMaxDepth = 40;       // deepest 8-wide tree the wide kernel's 56-entry stack takes with room to postpone triangle groups
constexpr size_t kWoopPadBytes = 64;   // private Woop copy: one terminator row + three zero rows past the caller's rows

}  // namespace b2rt

// One context per device.  Owns the private BVH copy, the persistent-thread work counter,
// scratch for the sort and the event pair used when the caller asks for kernel milliseconds.
struct b2rt_ctx {
    int device = 0;
    int smCount = 0, ccMajor = 0, ccMinor = 0;
    size_t l2Bytes = 0;
    std::string lastError;
    int64_t launches = 0;

    // private BVH copy (nodes re-laid: top levels first)
    float4* nodes = nullptr;   size_t nodeBytes = 0;
    float4* woop = nullptr;    size_t woopBytes = 0;
    int32_t* triIndex = nullptr; size_t triIndexBytes = 0;
    int topNodesAvail = 0;     // how many leading nodes are in breadth-first order
    bool hasBvh = false;

    // second private form: 8-wide compressed nodes + re-packed Woop rows (wide_bvh.h), built on first use
    float4* wideNodes = nullptr; float4* wideWoop = nullptr; int32_t* wideTriIndex = nullptr;
    int wideNumNodes = 0, wideNumTris = 0, wideMaxDepth = 0;
    size_t wideBytes = 0;       // the three arrays share one allocation (wideNodes is its start): one L2 access window covers them
    bool hasWide = false;
    bool wideGreedy = false;    // the form in place was built with B2RT_TUNE_GREEDY_WIDE

    b2rt_trace_config cfg = { -1, -1, -1, -1, -1, 0u };

    // Launch slots: every trace launch takes the next of kLaunchSlots work counters (and, for the any-hit kernels, the
    // next slice of the shared-ray lists), so traces issued on different streams may overlap -- up to kLaunchSlots in flight.
    static constexpr int kLaunchSlots = 8;
    int* warpCounter = nullptr;                 // kLaunchSlots work counters for persistent warps, 128 bytes apart
    unsigned launchSlot = 0;
    unsigned long long* statsDev = nullptr;     // 16 counters for the instrumented kernel + kTimelineWarps exit times
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // b2rt_trace_host: upload / trace / download pipeline
    static constexpr int kHostChunks = 32;
    static constexpr int kHostSlots = 2;        // host batches in flight (b2rt_trace_host_begin / _end)
    cudaStream_t sIn = nullptr, sRun = nullptr, sOut = nullptr;
    struct HostSlot {
        cudaEvent_t evIn[kHostChunks] = {}, evK0[kHostChunks] = {}, evK1[kHostChunks] = {}, evT0 = nullptr, evT1 = nullptr;
        void* stage = nullptr; size_t stageBytes = 0;       // device staging: rays, then results
        int chunks = 0;
        bool inFlight = false;
    } hostSlot[kHostSlots];
    int nextHostSlot = 0;

    // per (kernel, dynamic shared memory): resident CTAs per SM, asked once (the attribute is set at the same time)
    std::map<std::pair<const void*, size_t>, int> occupancy;
    size_t l2SetAside = 0;     // bytes of L2 set aside for persisting accesses by this context (0 = never asked)

    int sortMode = 0;          // b2rt_set_sort_mode: 0 = prefix sort + tie settling (falls back by itself), 1 = always the full LSD sort
    int sortFellBack = 0;      // the last b2rt_morton_sort went the long way

    // scratch (grown on demand, never shrunk)
    void* scratch = nullptr;  size_t scratchBytes = 0;
    int* shareList = nullptr; size_t shareListCtas = 0; // any-hit kernels: shared-ray lists, one slice per CTA (trace.cu)
    float* aabbDev = nullptr;                   // 6 floats
    long long* countDev = nullptr;              // hit counter: running sum + finished-block count (misc.cu)
    long long* countHost = nullptr;             // mapped pinned word the last block writes the total to
};

namespace b2rt {

int setError(b2rt_ctx* ctx, int code, const char* fmt, ...);
int ensureScratch(b2rt_ctx* ctx, size_t bytes);
extern std::string g_createError;

#define B2RT_CUDA(ctx, call)                                                                   \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return b2rt::setError((ctx), B2RT_ERR_CUDA, "%s failed: %s (%s:%d)", #call,        \
                                  cudaGetErrorString(e_), __FILE__, __LINE__);                 \
    } while (0)

#define B2RT_CHECK_LAUNCH(ctx, name)                                                           \
    do {                                                                                       \
        cudaError_t e_ = cudaGetLastError();                                                   \
        if (e_ != cudaSuccess)                                                                 \
            return b2rt::setError((ctx), B2RT_ERR_CUDA, "launch of %s failed: %s", name,       \
                                  cudaGetErrorString(e_));                                     \
        (ctx)->launches++;                                                                     \
    } while (0)

// The dynamic shared-memory limit of a kernel is a property of the (device, kernel) pair for the whole process, and the same
// kernel is launched with different sizes (the staged top of the tree depends on the BVH; several contexts may share a device):
// it is only ever RAISED, under a process-wide lock (api.cu).
int raiseSmemLimit(b2rt_ctx* ctx, const void* kern, size_t smem);

// cudaFuncSetAttribute + the occupancy query, once per (kernel, shared-memory size) instead of on every launch
template <class K>
int kernelOccupancy(b2rt_ctx* ctx, K kern, int block, size_t smem, int* blocksPerSm)
{
    const std::pair<const void*, size_t> key(reinterpret_cast<const void*>(kern), smem);
    auto it = ctx->occupancy.find(key);
    if (it == ctx->occupancy.end()) {
        int n = 0;
        { const int rc = raiseSmemLimit(ctx, reinterpret_cast<const void*>(kern), smem); if (rc) return rc; }
        B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, block, smem));
        if (n > 0 && !(ctx->cfg.tune & B2RT_TUNE_DEFAULT_CARVEOUT)) {
            // Shared memory and L1 share 228 KB per SM.  Left to itself the driver carved out room for more CTAs than the
            // registers allow (ncu: occupancy limit from shared memory 7 CTAs for a kernel that runs 4), i.e. ~64 KB of L1
            // lost to shared memory nobody uses, with the traversal's L1 hit rate at 68 %.  Ask for what the resident CTAs
            // need (their dynamic shared memory + 1 KB reserved per CTA) and leave the rest to L1.
            const size_t need = (size_t)n * (smem + 1024);
            int pct = (int)((need * 100 + 228 * 1024 - 1) / (228 * 1024));
            if (pct > 100) pct = 100;
            B2RT_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
            int n2 = 0;
            B2RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n2, kern, block, smem));
            if (n2 < n) B2RT_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutDefault));
        }
        it = ctx->occupancy.emplace(key, n).first;
    }
    *blocksPerSm = it->second;
    return B2RT_OK;
}

// kernels' launchers (defined in the .cu files)
int launchTrace(b2rt_ctx* ctx, const void* rays, void* results, int64_t numRays, unsigned flags,
                cudaStream_t stream);

int launchTraceWide(b2rt_ctx* ctx, const void* rays, void* results, int64_t numRays, unsigned flags, int variant,
                    cudaStream_t stream);
void freeWide(b2rt_ctx* ctx);
int ensureWide(b2rt_ctx* ctx);      // builds the wide form from the private Compact2 copy if it is not there yet
int uploadWide(b2rt_ctx* ctx, const void* nodes, size_t nodeBytes, const void* woop, size_t woopBytes, const void* tri,
               size_t triBytes, int numNodes, int numTris, int maxDepth, bool greedy, bool srcIsDevice);

int countWork(b2rt_ctx* ctx, const void* rays, int64_t numRays, int anyHit, b2rt_work_counters* out, cudaStream_t stream);

// device helpers
#ifdef __CUDACC__
__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }
__device__ __forceinline__ float rcp_approx(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// The Woop test's arithmetic, instruction for instruction what nvcc makes of kepler_dynamic_fetch.cu:336-358 in the reference's
// build (--use_fast_math: flush-to-zero FFMA / FMUL / FADD, MUFU.RCP; SASS of that file compiled for sm_100a, excerpt in
// profiles/r02_sass_woop_reference_vs_b2rt.md):
//   Oz    = fma(-oz, r0.z, fma(-oy, r0.y, fma(-ox, r0.x, r0.w)))
//   denom = fma( dz, r0.z, fma( dx, r0.x, dy * r0.y))          <- the PRODUCT is the y term: nvcc re-associates the sum
//   t     = Oz * rcp(denom)
//   Ox    = fma( oz, r1.z, fma( oy, r1.y, fma( ox, r1.x, r1.w))),  Dx = fma(dz, r1.z, fma(dx, r1.x, dy * r1.y)),  u = fma(t, Dx, Ox)
// (v likewise with row 2).  With the same operations in the same order t, u and v are bit-identical to the reference kernel's
// for the same ray and triangle; round 1 had the product on the x term and differed from it by up to 2.9e-5 relative on rays
// whose Oz cancels (secondary rays starting next to the plane they hit).
__device__ __forceinline__ float ffmaFtz(float a, float b, float c)
{
    float r;
    asm("fma.rn.ftz.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float fmulFtz(float a, float b)
{
    float r;
    asm("mul.rn.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float faddFtz(float a, float b)
{
    float r;
    asm("add.rn.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float woopDot(float dx, float dy, float dz, const float4 r) { return ffmaFtz(dz, r.z, ffmaFtz(dx, r.x, fmulFtz(dy, r.y))); }
__device__ __forceinline__ float woopPlane(float ox, float oy, float oz, const float4 r) { return ffmaFtz(oz, r.z, ffmaFtz(oy, r.y, ffmaFtz(ox, r.x, r.w))); }

// B2RT_TRACE_WATERTIGHT: the acceptance test of a triangle with an edge tolerance (the reference's is exact, u >= 0, v >= 0,
// u + v <= 1, kepler_dynamic_fetch.cu:349-364: a ray aimed at the edge two triangles share can fail both by one rounding and
// leak through a closed surface -- 2.8 % of the rays aimed at the diagonal of a split quad do, tests/test_gpu_watertight.py).
// u = r . P + w with P = o + t d the hit point and (r, w) a row of the float32 inverse CudaBVH.cc:361-380 stores: |r| is one
// over the triangle's edge length, so the row's rounding alone moves u by ulp x |r| |P| -- far more than an ulp of u's own
// terms, which nearly cancel (measured: v = -1.1e-6 from the stored rows for a point 5e-7 INSIDE the triangle).  The
// tolerance is therefore scaled by the row magnitudes:  eps_u = 4 ulp x (|r.x| |P.x| + |r.y| |P.y| + |r.z| |P.z| + |w|),  with
// |P.k| bounded by |o.k| + |t d.k| -- a position error of a few ulps of the coordinates, expressed in barycentric units.
// With it a ray on a shared edge hits BOTH triangles; which one is reported must then not depend on the order the traversal
// meets them in (that is: on the ray's warp-mates), so of two hits at exactly the same t the one with the lower triangle id
// wins, and boxes are culled against hitT widened by 2^-20 so that the other candidate of an exact tie is always still
// visited.  Results of a batch are then identical under any permutation of its rays.
constexpr float kWatertightEps = 4.0f / 16777216.0f;        // 4 ulps
struct WtRay { float ox, oy, oz, dx, dy, dz; };
template <bool WT>
__device__ __forceinline__ bool woopAccept(float t, float tmin, float hitT, float u, float v, const float4 r1, const float4 r2,
                                           const WtRay& q, int row, int hitRow, const int* __restrict__ triIndex)
{
    if (!WT) return t > tmin && t < hitT && u >= 0.0f && v >= 0.0f && faddFtz(u, v) <= 1.0f;
    const float hx = __fmaf_rn(fabsf(t), fabsf(q.dx), fabsf(q.ox)), hy = __fmaf_rn(fabsf(t), fabsf(q.dy), fabsf(q.oy)),
                hz = __fmaf_rn(fabsf(t), fabsf(q.dz), fabsf(q.oz));
    const float eu = kWatertightEps * __fmaf_rn(fabsf(r1.x), hx, __fmaf_rn(fabsf(r1.y), hy, __fmaf_rn(fabsf(r1.z), hz, fabsf(r1.w))));
    const float ev = kWatertightEps * __fmaf_rn(fabsf(r2.x), hx, __fmaf_rn(fabsf(r2.y), hy, __fmaf_rn(fabsf(r2.z), hz, fabsf(r2.w))));
    if (!(t > tmin && t <= hitT && u >= -eu && v >= -ev && u + v <= 1.0f + eu + ev)) return false;
    if (t < hitT) return true;
    return hitRow >= 0 && __ldg(triIndex + row) < __ldg(triIndex + hitRow);           // exact tie: the lower triangle id
}
__device__ __forceinline__ float watertightCull(float hitT) { return hitT * (1.0f + 1.0f / 1048576.0f); }
// ... and the BOX test has to be conservative too: shared vertices and edges are the extreme points of leaf boxes, and the slab
// distances fma(plane, idir, -orig * idir) (kepler_dynamic_fetch.cu:227-246) carry an ABSOLUTE error of an ulp of |orig * idir|
// (the product is rounded before the subtraction cancels it) plus the relative error of rcp.approx.  Watertight kernels accept
// a box when  tmax (1 + 2^-21) + slack >= tmin  with slack = 2^-21 max_axis |orig * idir|, computed once per ray.
__device__ __forceinline__ float watertightSlack(float oodx, float oody, float oodz)
{
    return (1.0f / 2097152.0f) * fmaxf(fmaxf(fabsf(oodx), fabsf(oody)), fabsf(oodz));
}
__device__ __forceinline__ bool watertightOverlap(float cmin, float cmax, float slack)
{
    return __fmaf_rn(cmax, 1.0f + 1.0f / 2097152.0f, slack) >= cmin;
}

// shared-memory loads through a 32-bit shared-window address (read-only data staged before a barrier)
__device__ __forceinline__ float4 lds128(unsigned addr)
{
    float4 v;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ int2 lds64i(unsigned addr)
{
    int2 v;
    asm("ld.shared.v2.s32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ unsigned long long globalTimerNs()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// sm_100a three-input float min/max (SASS FMNMX3)
__device__ __forceinline__ float fmin3(float a, float b, float c)
{
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float fmax3(float a, float b, float c)
{
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
#endif

}  // namespace b2rt
