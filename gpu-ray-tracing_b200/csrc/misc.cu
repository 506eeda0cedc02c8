// misc.cu -- hit counting (the numerator of the reported Mrays/s) and pixel reconstruction.
//
// Replaces cuda/RendererKernels.cu:112-162 (countHitsKernel; caller Renderer::getTotalNumRays,
// Renderer.cc:221-238) and :60-108 (reconstructKernel) with their launchers :166-212.
#include "common.cuh"

namespace b2rt {

// state[0] = running sum, state[1] = finished blocks.  The last block to finish hands the total to the host through a
// mapped pinned word and leaves both words zero for the next call: no memset before and no copy after the launch.
__global__ void __launch_bounds__(256) countHitsKernel(const int4* __restrict__ results, long long n,
                                                       unsigned long long* state, unsigned long long* hostOut)
{
    unsigned c = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        c += (__ldg(reinterpret_cast<const int*>(results + i)) >= 0) ? 1u : 0u;
    c = __reduce_add_sync(0xffffffffu, c);
    __shared__ unsigned s[8];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) t += s[w];
        if (t) atomicAdd(&state[0], t);
        __threadfence();
        if (atomicAdd(&state[1], 1ull) == (unsigned long long)gridDim.x - 1) {
            __threadfence();
            *hostOut = atomicExch(&state[0], 0ull);
            state[1] = 0;
            __threadfence_system();
        }
    }
}

// Vec4f::fromABGR / toABGR -- framework/base/Math.cc:34-54
__device__ __forceinline__ float4 fromABGR(unsigned abgr)
{
    return make_float4((float)(abgr & 0xFF) * (1.0f / 255.0f), (float)((abgr >> 8) & 0xFF) * (1.0f / 255.0f),
                       (float)((abgr >> 16) & 0xFF) * (1.0f / 255.0f), (float)(abgr >> 24) * (1.0f / 255.0f));
}
__device__ __forceinline__ unsigned chanToU8(float x)
{
    const float cl = fminf(fmaxf(x, 0.0f), 1.0f);
    return (unsigned)((((unsigned long long)(cl * 72057594037927936.0f) * 255ull) >> 55) + 1ull) >> 1;
}
__device__ __forceinline__ unsigned toABGR(float4 c)
{
    return chanToU8(c.x) | (chanToU8(c.y) << 8) | (chanToU8(c.z) << 16) | (chanToU8(c.w) << 24);
}

struct ReconstructParams {
    int numRaysPerPrimary, firstPrimary, numPrimary, mode;      // mode: 0 primary, 1 AO, 2 diffuse
    const int32_t* primarySlotToID;
    const int4* primaryResults;
    const int32_t* batchIDToSlot;
    const int4* batchResults;
    const unsigned* triMaterialColor;
    const unsigned* triShadedColor;
    unsigned* pixels;
};

__global__ void __launch_bounds__(256) reconstructKernel(const ReconstructParams p)
{
    const int taskIdx = blockIdx.x * blockDim.x + threadIdx.x;
    if (taskIdx >= p.numPrimary) return;
    const bool isPrimary = p.mode == 0, isAO = p.mode == 1, isDiffuse = p.mode == 2;
    const int primarySlot = p.firstPrimary + taskIdx;
    const int primaryID = p.primarySlotToID[primarySlot];
    const int primaryTri = p.primaryResults[primarySlot].x;
    const int32_t* batchSlots = p.batchIDToSlot + (isPrimary ? primaryID : taskIdx * p.numRaysPerPrimary);
    const float4 bg = make_float4(0.2f, 0.4f, 0.8f, 1.0f);
    float4 color = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = 0; i < p.numRaysPerPrimary; i++) {
        const int tri = p.batchResults[batchSlots[i]].x;
        float4 add;
        if (tri == -1) add = isPrimary ? bg : make_float4(1.f, 1.f, 1.f, 1.f);
        else add = isAO ? make_float4(0.f, 0.f, 0.f, 1.f) : fromABGR(p.triShadedColor[tri]);
        color.x += add.x; color.y += add.y; color.z += add.z; color.w += add.w;
    }
    const float s = 1.0f / (float)p.numRaysPerPrimary;
    color.x *= s; color.y *= s; color.z *= s; color.w *= s;
    if (isAO && primaryTri == -1) color = bg;
    if (isDiffuse) {
        const float4 m = (primaryTri == -1) ? bg : fromABGR(p.triMaterialColor[primaryTri]);
        color.x *= m.x; color.y *= m.y; color.z *= m.z; color.w *= m.w;
    }
    p.pixels[primaryID] = toABGR(color);
}

}  // namespace b2rt

using namespace b2rt;

extern "C" {

int b2rt_count_hits(b2rt_ctx* ctx, const void* results, int64_t n, int64_t* hits, void* stream)
{
    if (!ctx || !hits) return B2RT_ERR_ARG;
    *hits = 0;
    if (n < 0 || (n && !results)) return setError(ctx, B2RT_ERR_ARG, "b2rt_count_hits: bad arguments");
    if (n == 0) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    unsigned long long* state = reinterpret_cast<unsigned long long*>(ctx->countDev);
    long long grid = (n + 255) / 256;
    const long long cap = (long long)ctx->smCount * 8;
    if (grid > cap) grid = cap;
    countHitsKernel<<<(unsigned)grid, 256, 0, s>>>(static_cast<const int4*>(results), n, state,
                                                   reinterpret_cast<unsigned long long*>(ctx->countHost));
    B2RT_CHECK_LAUNCH(ctx, "countHitsKernel");
    B2RT_CUDA(ctx, cudaStreamSynchronize(s));
    const unsigned long long h = *reinterpret_cast<volatile unsigned long long*>(ctx->countHost);
    *hits = (int64_t)h;
    return B2RT_OK;
}

int b2rt_reconstruct(b2rt_ctx* ctx, int numRaysPerPrimary, int firstPrimary, int numPrimary, int mode,
                     const int32_t* primarySlotToID, const void* primaryResults, const int32_t* batchIDToSlot,
                     const void* batchResults, const uint32_t* triMaterialColor, const uint32_t* triShadedColor,
                     uint32_t* pixels, void* stream)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (numPrimary < 0 || numRaysPerPrimary < 1 || mode < 0 || mode > 2)
        return setError(ctx, B2RT_ERR_ARG, "b2rt_reconstruct: bad arguments");
    if (numPrimary == 0) return B2RT_OK;
    if (!primarySlotToID || !primaryResults || !batchIDToSlot || !batchResults || !pixels ||
        (mode != 1 && !triShadedColor) || (mode == 2 && !triMaterialColor))
        return setError(ctx, B2RT_ERR_ARG, "b2rt_reconstruct: NULL argument");
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    ReconstructParams p;
    p.numRaysPerPrimary = numRaysPerPrimary; p.firstPrimary = firstPrimary; p.numPrimary = numPrimary; p.mode = mode;
    p.primarySlotToID = primarySlotToID; p.primaryResults = static_cast<const int4*>(primaryResults);
    p.batchIDToSlot = batchIDToSlot; p.batchResults = static_cast<const int4*>(batchResults);
    p.triMaterialColor = triMaterialColor; p.triShadedColor = triShadedColor; p.pixels = pixels;
    reconstructKernel<<<(numPrimary + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
    B2RT_CHECK_LAUNCH(ctx, "reconstructKernel");
    return B2RT_OK;
}

}  // extern "C"
