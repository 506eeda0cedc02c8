// refabi.cpp -- libb2rt_refabi.so: the reference's C-linkage kernel launchers (include/b2rt_refabi.h)
// forwarded to the C ABI of libb2rt.so.  No kernels here; each function cites what it replaces.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include <map>
#include <mutex>

#include "../../../include/b2rt.h"
#include "../../../include/b2rt_refabi.h"

namespace {

std::mutex g_lock;
std::map<int, b2rt_ctx*> g_ctx;        // one per CUDA device, like the reference's per-process texture bindings

// FW::fail / cutilSafeCall convention: message, then exit(-1)  (Defs.cc:41-54, cutil_inline_runtime.h:33-42)
[[noreturn]] void die(const char* what, const b2rt_ctx* ctx)
{
    fprintf(stderr, "%s: %s\n", what, b2rt_last_error(ctx));
    exit(-1);
}

b2rt_ctx* context()
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { fprintf(stderr, "b2rt_refabi: cudaGetDevice failed\n"); exit(-1); }
    std::lock_guard<std::mutex> guard(g_lock);
    auto it = g_ctx.find(dev);
    if (it != g_ctx.end()) return it->second;
    b2rt_ctx* ctx = nullptr;
    if (b2rt_create(dev, &ctx) != B2RT_OK) die("b2rt_create", nullptr);
    g_ctx[dev] = ctx;
    return ctx;
}

inline void check(int rc, const char* what, b2rt_ctx* ctx) { if (rc != B2RT_OK) die(what, ctx); }

}  // namespace

extern "C" {

// kepler_dynamic_fetch.cu:417-424 (cudaBindTexture x3) -> private re-laid copy inside the context
void bind_CudaBVHTexture(void* nodes, int64_t nodeBytes, void* woop, int64_t woopBytes, int* triIndices, int64_t triIndexBytes)
{
    b2rt_ctx* ctx = context();
    check(b2rt_set_bvh(ctx, nodes, (size_t)nodeBytes, woop, (size_t)woopBytes, triIndices, (size_t)triIndexBytes,
                       B2RT_LAYOUT_COMPACT2, 1), "bind_CudaBVHTexture", ctx);
}

// kepler_dynamic_fetch.cu:426-430: the context keeps its copy until the next bind or process exit
void unbind_CudaBVHTexture(void) {}

// kepler_dynamic_fetch.cu:432-474
float launch_tracingKernel(int32_t, int*, int numRays, bool anyHit, void* rays, void* results, void*, void*, void*, void*, void*,
                           void*, void*, int*)
{
    b2rt_ctx* ctx = context();
    float ms = 0.0f;
    check(b2rt_trace(ctx, rays, results, numRays, anyHit ? B2RT_TRACE_ANY_HIT : 0u, nullptr, &ms), "launch_tracingKernel", ctx);
    return ms;
}

// CudaTracerKernels.hh:52
void copy_tracing_results(void* hostResults, void* devResults, int32_t n)
{
    if (n > 0 && cudaMemcpy(hostResults, devResults, (size_t)n * 16, cudaMemcpyDeviceToHost) != cudaSuccess) {
        fprintf(stderr, "copy_tracing_results: cudaMemcpy failed\n");
        exit(-1);
    }
}

// RayGenKernels.cu:295-310 (synchronous: ends with cudaDeviceSynchronize)
void launch_rayGenPrimaryKernel(int32_t, b2rt_ref_RayGenPrimaryInput* in)
{
    b2rt_ctx* ctx = context();
    check(b2rt_raygen_primary(ctx, in->origin, in->nscreenToWorld, in->w, in->h, in->maxDist, in->indexToPixel, in->rays,
                              in->idToSlot, in->slotToID, nullptr), "launch_rayGenPrimaryKernel", ctx);
    cudaDeviceSynchronize();
}

// RayGenKernels.cu:312-330
void launch_rayGenAOKernel(int32_t, b2rt_ref_RayGenAOInput* in)
{
    b2rt_ctx* ctx = context();
    check(b2rt_raygen_ao(ctx, in->firstInputSlot, in->numInputRays, in->numSamples, in->maxDist, in->randomSeed, in->inRays,
                         in->inResults, static_cast<const float*>(in->normals), in->outRays, in->outIDToSlot, in->outSlotToID,
                         nullptr), "launch_rayGenAOKernel", ctx);
    cudaDeviceSynchronize();
}

// RayBufferKernels.cu:204-221
void launch_findAABBKernel(int32_t, const int*, b2rt_ref_FindAABBInput* in, b2rt_ref_FindAABBOutput* out)
{
    b2rt_ctx* ctx = context();
    float box[6];
    check(b2rt_find_aabb(ctx, in->inRays, in->numRays, box, nullptr), "launch_findAABBKernel", ctx);
    for (int k = 0; k < 3; k++) { out->aabbLo[k] = box[k]; out->aabbHi[k] = box[3 + k]; }
}

// RayBufferKernels.cu:223-238
void launch_genMortonKeysKernel(int32_t, const int*, b2rt_ref_GenMortonKeysInput* in)
{
    b2rt_ctx* ctx = context();
    const float box[6] = { in->aabbLo[0], in->aabbLo[1], in->aabbLo[2], in->aabbHi[0], in->aabbHi[1], in->aabbHi[2] };
    check(b2rt_morton_keys(ctx, in->inRays, in->numRays, box, in->outKeys, nullptr), "launch_genMortonKeysKernel", ctx);
}

// RayBufferKernels.cu:240-255
void launch_reorderRaysKernel(int32_t, b2rt_ref_ReorderRaysInput* in)
{
    b2rt_ctx* ctx = context();
    check(b2rt_reorder_rays(ctx, in->inKeys, in->inRays, in->inSlotToID, in->outRays, in->outIDToSlot, in->outSlotToID, in->numRays,
                            nullptr), "launch_reorderRaysKernel", ctx);
    cudaDeviceSynchronize();
}

// RendererKernels.cu:164-187
void launch_reconstructKernel(int32_t, b2rt_ref_ReconstructInput* in)
{
    b2rt_ctx* ctx = context();
    const int mode = in->isPrimary ? 0 : (in->isAO ? 1 : 2);
    check(b2rt_reconstruct(ctx, in->numRaysPerPrimary, in->firstPrimary, in->numPrimary, mode, in->primarySlotToID, in->primaryResults,
                           in->batchIDToSlot, in->batchResults, in->triMaterialColor, in->triShadedColor, in->pixels, nullptr),
          "launch_reconstructKernel", ctx);
    cudaDeviceSynchronize();
}

// RendererKernels.cu:189-212
int32_t launch_countHitsKernel(int32_t, const int*, b2rt_ref_CountHitsInput* in)
{
    b2rt_ctx* ctx = context();
    int64_t hits = 0;
    check(b2rt_count_hits(ctx, in->rayResults, in->numRays, &hits, nullptr), "launch_countHitsKernel", ctx);
    return (int32_t)hits;
}

}  // extern "C"
