// b2rt_fw.hpp -- header-only C++ mirror of the reference's classes for the ray-casting path, over
// the C ABI of libb2rt.so (include/b2rt.h).  Same names, argument meaning and error behaviour as
//   FW::CudaBVH     src/rt/cuda/CudaBVH.hh:86-137
//   FW::RayBuffer   src/rt/ray/RayBuffer.hh:37-97
//   FW::RayGen      src/rt/ray/RayGen.hh:41-69
//   FW::CudaTracer  src/rt/cuda/CudaTracer.hh:39-69
// so code written against the reference compiles against this header with `namespace FW = b2rt_fw`.
// Device memory is plain cudaMalloc, as in the reference (RayBuffer.cc:53-81, CudaBVH.cc:350-356).
// Errors: FW::fail() prints and exits in the reference (Defs.cc:41-54); here fail() throws
// std::runtime_error with the same message so tests can observe it.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <stdexcept>
#include <string>
#include <vector>

#include "b2rt.h"

namespace b2rt_fw {

typedef int32_t S32;
typedef uint32_t U32;
typedef float F32;

struct Vec3f { F32 x, y, z; };
struct Ray { Vec3f origin; F32 tmin; Vec3f direction; F32 tmax; };                    // rt/Util.hh:64-73
struct RayResult { S32 id; F32 t; S32 padA, padB; bool hit() const { return id != -1; } };   // rt/Util.hh:79-89
enum { RAY_NO_HIT = -1 };
enum BVHLayout { BVHLayout_Compact2 = B2RT_LAYOUT_COMPACT2 };                          // CudaTracerKernels.hh:117-127
static_assert(sizeof(Ray) == 32 && sizeof(RayResult) == 16, "ABI");

inline void fail(const std::string& msg) { throw std::runtime_error(msg); }
inline void cudaCheck(cudaError_t e, const char* what)
{
    if (e != cudaSuccess) fail(std::string(what) + ": " + cudaGetErrorString(e));
}

// One context per device, created on first use (replaces CudaModule::staticInit + CudaTracer::s_inited).
inline b2rt_ctx* context()
{
    static b2rt_ctx* ctx = nullptr;
    if (!ctx) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (b2rt_create(dev, &ctx) != B2RT_OK) fail(b2rt_last_error(nullptr));
    }
    return ctx;
}
inline void check(int rc) { if (rc != B2RT_OK) fail(b2rt_last_error(context())); }

// ---------------------------------------------------------------------------------------------
class CudaBVH {
public:
    // from a host BVH built by b2rt_bvh_build / loaded by b2rt_bvh_load  (CudaBVH(BVH&, BVHLayout), CudaBVH.cc:66-70)
    explicit CudaBVH(const b2rt_host_bvh* bvh, BVHLayout layout = BVHLayout_Compact2) : m_layout(layout)
    {
        const void *n, *w; const int32_t* t;
        check(b2rt_bvh_arrays(bvh, &n, &m_nodeBytes, &w, &m_woopBytes, &t, &m_triBytes));
        upload(n, w, t);
    }
    // from raw Compact2 arrays (CudaBVH(InputStream&), CudaBVH.cc:79-97)
    CudaBVH(const void* nodes, size_t nodeBytes, const void* woop, size_t woopBytes, const int32_t* tri, size_t triBytes,
            BVHLayout layout = BVHLayout_Compact2)
        : m_layout(layout), m_nodeBytes(nodeBytes), m_woopBytes(woopBytes), m_triBytes(triBytes) { upload(nodes, woop, tri); }
    ~CudaBVH() { cudaFree(m_nodes_dev); cudaFree(m_triWoop_dev); cudaFree(m_triIndex_dev); }
    CudaBVH(const CudaBVH&) = delete;
    CudaBVH& operator=(const CudaBVH&) = delete;

    BVHLayout getLayout() const { return m_layout; }
    void* getNodeBuffer_dev() const { return m_nodes_dev; }
    void* getTriWoopBuffer_dev() const { return m_triWoop_dev; }
    S32* getTriIndexBuffer_dev() const { return m_triIndex_dev; }
    size_t getNodeBytes() const { return m_nodeBytes; }
    size_t getTriWoopBytes() const { return m_woopBytes; }
    size_t getTriIndexBytes() const { return m_triBytes; }

private:
    void upload(const void* n, const void* w, const int32_t* t)
    {
        cudaCheck(cudaMalloc(&m_nodes_dev, m_nodeBytes), "cudaMalloc nodes");
        cudaCheck(cudaMalloc(&m_triWoop_dev, m_woopBytes), "cudaMalloc woop");
        cudaCheck(cudaMalloc((void**)&m_triIndex_dev, m_triBytes), "cudaMalloc triIndex");
        cudaCheck(cudaMemcpy(m_nodes_dev, n, m_nodeBytes, cudaMemcpyHostToDevice), "copy nodes");
        cudaCheck(cudaMemcpy(m_triWoop_dev, w, m_woopBytes, cudaMemcpyHostToDevice), "copy woop");
        cudaCheck(cudaMemcpy(m_triIndex_dev, t, m_triBytes, cudaMemcpyHostToDevice), "copy triIndex");
    }
    BVHLayout m_layout;
    size_t m_nodeBytes = 0, m_woopBytes = 0, m_triBytes = 0;
    void* m_nodes_dev = nullptr; void* m_triWoop_dev = nullptr; S32* m_triIndex_dev = nullptr;
};

// ---------------------------------------------------------------------------------------------
class RayBuffer {
public:
    explicit RayBuffer(S32 n = 0, bool closestHit = true) : m_needClosestHit(closestHit) { resize(n); }
    ~RayBuffer() { release(); }
    RayBuffer(const RayBuffer&) = delete;
    RayBuffer& operator=(const RayBuffer&) = delete;

    S32 getSize() const { return m_size; }
    // RayBuffer.cc:53-81: shrinking is logical; growing frees and reallocates (contents lost)
    void resize(S32 n)
    {
        if (n < 0) fail("RayBuffer::resize: negative size");
        if (n <= m_capacity) { m_size = n; return; }
        release();
        m_size = m_capacity = n;
        cudaCheck(cudaMalloc((void**)&m_rays_dev, (size_t)n * sizeof(Ray)), "cudaMalloc rays");
        cudaCheck(cudaMalloc((void**)&m_results_dev, (size_t)n * sizeof(RayResult)), "cudaMalloc results");
        cudaCheck(cudaMalloc((void**)&m_IDToSlot_dev, (size_t)n * 4), "cudaMalloc idToSlot");
        cudaCheck(cudaMalloc((void**)&m_slotToID_dev, (size_t)n * 4), "cudaMalloc slotToID");
    }
    void setNeedClosestHit(bool c) { m_needClosestHit = c; }
    bool getNeedClosestHit() const { return m_needClosestHit; }

    Ray* getRayBuffer_dev() const { return m_rays_dev; }
    RayResult* getResultBuffer_dev() const { return m_results_dev; }
    S32* getIDToSlotBuffer_dev() const { return m_IDToSlot_dev; }
    S32* getSlotToIDBuffer_dev() const { return m_slotToID_dev; }

    // host <-> device helpers (setRay / getResultForSlot one by one in the reference, RayBuffer.hh:46-60)
    void setRays(const std::vector<Ray>& rays)
    {
        resize((S32)rays.size());
        std::vector<S32> ident(rays.size());
        for (size_t i = 0; i < ident.size(); i++) ident[i] = (S32)i;
        if (rays.empty()) return;
        cudaCheck(cudaMemcpy(m_rays_dev, rays.data(), rays.size() * sizeof(Ray), cudaMemcpyHostToDevice), "copy rays");
        cudaCheck(cudaMemcpy(m_IDToSlot_dev, ident.data(), ident.size() * 4, cudaMemcpyHostToDevice), "copy ids");
        cudaCheck(cudaMemcpy(m_slotToID_dev, ident.data(), ident.size() * 4, cudaMemcpyHostToDevice), "copy ids");
    }
    std::vector<Ray> getRays() const { return download<Ray>(m_rays_dev); }
    std::vector<RayResult> getResults() const { return download<RayResult>(m_results_dev); }
    std::vector<S32> getSlotToID() const { return download<S32>(m_slotToID_dev); }
    std::vector<S32> getIDToSlot() const { return download<S32>(m_IDToSlot_dev); }

    // RayBuffer::mortonSort (RayBuffer.cc:256-324), entirely on the device
    void mortonSort()
    {
        float box[6];
        check(b2rt_morton_sort(context(), m_rays_dev, m_IDToSlot_dev, m_slotToID_dev, m_size, box, nullptr));
        m_aabbLo = Vec3f{ box[0], box[1], box[2] };
        m_aabbHi = Vec3f{ box[3], box[4], box[5] };
    }
    void getRayAABB(Vec3f& lo, Vec3f& hi) const { lo = m_aabbLo; hi = m_aabbHi; }

private:
    template <class T> std::vector<T> download(const T* dev) const
    {
        std::vector<T> v((size_t)m_size);
        if (m_size) cudaCheck(cudaMemcpy(v.data(), dev, v.size() * sizeof(T), cudaMemcpyDeviceToHost), "download");
        return v;
    }
    void release()
    {
        cudaFree(m_rays_dev); cudaFree(m_results_dev); cudaFree(m_IDToSlot_dev); cudaFree(m_slotToID_dev);
        m_rays_dev = nullptr; m_results_dev = nullptr; m_IDToSlot_dev = nullptr; m_slotToID_dev = nullptr;
        m_size = m_capacity = 0;
    }
    S32 m_size = 0, m_capacity = 0;
    bool m_needClosestHit;
    Ray* m_rays_dev = nullptr; RayResult* m_results_dev = nullptr; S32* m_IDToSlot_dev = nullptr; S32* m_slotToID_dev = nullptr;
    Vec3f m_aabbLo{ 0, 0, 0 }, m_aabbHi{ 0, 0, 0 };
};

// ---------------------------------------------------------------------------------------------
class RayGen {
public:
    explicit RayGen(S32 maxBatchSize = 8 * 1024 * 1024) : m_maxBatchSize(maxBatchSize) { srand(1); }
    ~RayGen() { cudaFree(m_indexToPixel_dev); cudaFree(m_normals_dev); }

    // RayGen::primary (RayGen.cc:50-73): nscreenToWorld column-major like FW::Mat4f
    void primary(RayBuffer& orays, const Vec3f& origin, const float nscreenToWorld[16], S32 w, S32 h, F32 maxDist)
    {
        if (w != m_w || h != m_h) {                                   // PixelTable::setSize (PixelTable.cc:55-68)
            std::vector<S32> lut((size_t)w * h);
            check(b2rt_pixel_table(w, h, lut.data(), nullptr));
            cudaFree(m_indexToPixel_dev);
            cudaCheck(cudaMalloc((void**)&m_indexToPixel_dev, lut.size() * 4), "cudaMalloc pixel table");
            cudaCheck(cudaMemcpy(m_indexToPixel_dev, lut.data(), lut.size() * 4, cudaMemcpyHostToDevice), "copy pixel table");
            m_w = w; m_h = h;
        }
        orays.resize(w * h);
        orays.setNeedClosestHit(true);
        const float o[3] = { origin.x, origin.y, origin.z };
        check(b2rt_raygen_primary(context(), o, nscreenToWorld, w, h, maxDist, m_indexToPixel_dev, orays.getRayBuffer_dev(),
                                  orays.getIDToSlotBuffer_dev(), orays.getSlotToIDBuffer_dev(), nullptr));
        cudaCheck(cudaDeviceSynchronize(), "rayGenPrimary");        // RayGenKernels.cu:309
    }

    // Scene::getTriNormalBuffer: uploaded once here, per batch in the reference (RayGen.cc:87-91)
    void setTriNormals(const float* normals, S32 numTris)
    {
        cudaFree(m_normals_dev);
        cudaCheck(cudaMalloc((void**)&m_normals_dev, (size_t)numTris * 12), "cudaMalloc normals");
        cudaCheck(cudaMemcpy(m_normals_dev, normals, (size_t)numTris * 12, cudaMemcpyHostToDevice), "copy normals");
    }

    // RayGen::ao (RayGen.cc:77-120); AO and Diffuse share it (Renderer.cc:268-277).  As in the reference the
    // randomSeed argument is ignored and the seed is libc rand() (RayGen.cc:106).
    bool ao(RayBuffer& orays, RayBuffer& irays, int numSamples, float maxDist, bool& newBatch, U32 /*randomSeed*/ = 0)
    {
        S32 lo, hi;
        if (!batching(irays.getSize(), numSamples, m_aoStartIdx, newBatch, lo, hi)) return false;
        orays.resize((hi - lo) * numSamples);
        orays.setNeedClosestHit(false);
        m_lastSeed = (U32)rand();
        check(b2rt_raygen_ao(context(), lo, hi - lo, numSamples, maxDist, m_lastSeed, irays.getRayBuffer_dev(),
                             irays.getResultBuffer_dev(), m_normals_dev, orays.getRayBuffer_dev(), orays.getIDToSlotBuffer_dev(),
                             orays.getSlotToIDBuffer_dev(), nullptr));
        cudaCheck(cudaDeviceSynchronize(), "rayGenAO");
        return true;
    }
    U32 lastSeed() const { return m_lastSeed; }

private:
    // RayGen::batching (RayGen.cc:124-142)
    bool batching(S32 numInputRays, S32 numSamples, S32& startIdx, bool& newBatch, S32& lo, S32& hi)
    {
        if (newBatch) { newBatch = false; startIdx = 0; }
        if (startIdx == numInputRays) return false;
        lo = startIdx;
        hi = lo + m_maxBatchSize / numSamples;
        if (hi > numInputRays) hi = numInputRays;
        startIdx = hi;
        return true;
    }
    S32 m_maxBatchSize, m_aoStartIdx = 0, m_w = 0, m_h = 0;
    U32 m_lastSeed = 0;
    S32* m_indexToPixel_dev = nullptr;
    float* m_normals_dev = nullptr;
};

// ---------------------------------------------------------------------------------------------
class CudaTracer {
public:
    CudaTracer() { initKernel(); }
    void initKernel() {}                                              // CudaTracer.cc:84-115: Compact2, persistent threads
    BVHLayout getDesiredBVHLayout() const { return BVHLayout_Compact2; }
    // CudaTracer.hh:49 + the texture bind at CudaTracer.cc:142-146
    void setBVH(CudaBVH* bvh)
    {
        m_bvh = bvh;
        if (bvh && bvh->getLayout() == getDesiredBVHLayout())
            check(b2rt_set_bvh(context(), bvh->getNodeBuffer_dev(), bvh->getNodeBytes(), bvh->getTriWoopBuffer_dev(),
                               bvh->getTriWoopBytes(), bvh->getTriIndexBuffer_dev(), bvh->getTriIndexBytes(), bvh->getLayout(), 1));
    }
    // CudaTracer::traceBatch (CudaTracer.cc:119-178): kernel time in MILLISECONDS
    F32 traceBatch(RayBuffer& rays)
    {
        if (!rays.getSize()) return 0.0f;
        if (!m_bvh) fail("CudaTracer: No BVH!");
        if (m_bvh->getLayout() != getDesiredBVHLayout()) fail("CudaTracer: Incorrect BVH layout!");
        float ms = 0.0f;
        unsigned flags = rays.getNeedClosestHit() ? 0u : B2RT_TRACE_ANY_HIT;
        if (writeBarycentrics) flags |= B2RT_TRACE_WRITE_BARY;
        check(b2rt_trace(context(), rays.getRayBuffer_dev(), rays.getResultBuffer_dev(), rays.getSize(), flags, nullptr, &ms));
        return ms;
    }
    bool second_fetch = false;                                        // CudaTracer.hh:52 (dump tooling flag, unused)
    bool writeBarycentrics = false;                                   // superset: u,v into RayResult.padA/padB

private:
    CudaBVH* m_bvh = nullptr;
};

// Renderer::getTotalNumRays' building block (Renderer.cc:221-238, RendererKernels.cu:112-162)
inline S32 countHits(const RayBuffer& rays)
{
    int64_t hits = 0;
    check(b2rt_count_hits(context(), rays.getResultBuffer_dev(), rays.getSize(), &hits, nullptr));
    return (S32)hits;
}

}  // namespace b2rt_fw
