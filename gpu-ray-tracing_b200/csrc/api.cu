// api.cu -- context, BVH upload and the trace entry points of the C ABI (include/b2rt.h).
#include "common.cuh"
#include "wide_bvh.h"

#include <stdarg.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <queue>
#include <vector>

namespace b2rt {

std::string g_createError;

int setError(b2rt_ctx* ctx, int code, const char* fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->lastError = buf; else g_createError = buf;
    return code;
}

int raiseSmemLimit(b2rt_ctx* ctx, const void* kern, size_t smem)
{
    static std::mutex lock;
    static std::map<std::pair<int, const void*>, size_t> limit;
    std::lock_guard<std::mutex> guard(lock);
    size_t& cur = limit[std::make_pair(ctx->device, kern)];
    if (smem > cur) {
        B2RT_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cur = smem;
    }
    return B2RT_OK;
}

int ensureScratch(b2rt_ctx* ctx, size_t bytes)
{
    if (bytes <= ctx->scratchBytes) return B2RT_OK;
    if (ctx->scratch) { cudaFree(ctx->scratch); ctx->scratch = nullptr; ctx->scratchBytes = 0; }
    B2RT_CUDA(ctx, cudaMalloc(&ctx->scratch, bytes));
    ctx->scratchBytes = bytes;
    return B2RT_OK;
}

static int ensureStage(b2rt_ctx* ctx, b2rt_ctx::HostSlot& slot, size_t bytes)
{
    if (bytes <= slot.stageBytes) return B2RT_OK;
    if (slot.stage) { cudaFree(slot.stage); slot.stage = nullptr; slot.stageBytes = 0; }
    B2RT_CUDA(ctx, cudaMalloc(&slot.stage, bytes));
    slot.stageBytes = bytes;
    return B2RT_OK;
}

// How many leading nodes of the private copy are ordered "most likely visited first".
constexpr int kMaxTopNodes = 4096;

// Re-lay the node array: the kMaxTopNodes nodes with the largest boxes (a greedy descent from the
// root by surface area, i.e. by SAH visit probability) come first, the rest keep their order.
// Child references (float4 indices, CudaBVH.cc:303) are rewritten; leaf references (~woop row,
// CudaBVH.cc:312) are untouched, so hit indices still address the caller's Woop / triIndex rows.
static void relayoutNodes(const std::vector<float>& in, std::vector<float>& out, int& topCount)
{
    const int n = (int)(in.size() / 16);
    out.resize(in.size());
    topCount = 0;
    if (n == 0) return;
    auto childArea = [&](int node, int c) -> float {
        const float* p = &in[(size_t)node * 16];
        float dx, dy, dz;
        if (c == 0) { dx = p[1] - p[0]; dy = p[3] - p[2]; dz = p[9] - p[8]; }
        else        { dx = p[5] - p[4]; dy = p[7] - p[6]; dz = p[11] - p[10]; }
        return dx * dy + dy * dz + dz * dx;
    };
    std::vector<int> order; order.reserve(n);
    std::vector<char> placed(n, 0);
    typedef std::pair<float, int> Item;
    std::priority_queue<Item> pq;
    pq.push(Item(3.0e38f, 0));
    const int want = std::min(n, kMaxTopNodes);
    while (!pq.empty() && (int)order.size() < want) {
        const int node = pq.top().second; pq.pop();
        if (placed[node]) continue;
        placed[node] = 1; order.push_back(node);
        int32_t ch[2]; memcpy(ch, &in[(size_t)node * 16 + 12], 8);
        for (int c = 0; c < 2; c++)
            if (ch[c] >= 0 && ch[c] / 4 < n) pq.push(Item(childArea(node, c), ch[c] / 4));
    }
    topCount = (int)order.size();
    for (int i = 0; i < n; i++) if (!placed[i]) order.push_back(i);
    std::vector<int> newIndex(n);
    for (int i = 0; i < n; i++) newIndex[order[i]] = i;
    for (int i = 0; i < n; i++) {
        const float* src = &in[(size_t)order[i] * 16];
        float* dst = &out[(size_t)i * 16];
        memcpy(dst, src, 64);
        int32_t ch[4]; memcpy(ch, src + 12, 16);
        for (int c = 0; c < 2; c++)
            if (ch[c] >= 0 && ch[c] / 4 < n) ch[c] = newIndex[ch[c] / 4] * 4;
        memcpy(dst + 12, ch, 16);
    }
}

void freeWide(b2rt_ctx* ctx)
{
    cudaFree(ctx->wideNodes);                   // one allocation: nodes, Woop rows, triangle ids
    ctx->wideNodes = nullptr; ctx->wideWoop = nullptr; ctx->wideTriIndex = nullptr;
    ctx->wideNumNodes = ctx->wideNumTris = ctx->wideMaxDepth = 0;
    ctx->wideBytes = 0;
    ctx->hasWide = false;
}

// Collapse the private Compact2 copy (node order is irrelevant to the collapse; the root is node 0 in both orders)
// into the 8-wide compressed form.  One-off host work per BVH: D2H of the three arrays, collapse, H2D.
int ensureWide(b2rt_ctx* ctx)
{
    const bool greedy = (ctx->cfg.tune & B2RT_TUNE_GREEDY_WIDE) != 0;
    if (ctx->hasWide && ctx->wideGreedy == greedy) return B2RT_OK;
    if (!ctx->hasBvh) return setError(ctx, B2RT_ERR_NO_BVH, "CudaTracer: No BVH!");
    std::vector<float> nodes(ctx->nodeBytes / 4), woop(ctx->woopBytes / 4);
    std::vector<int32_t> tri(ctx->triIndexBytes / 4);
    B2RT_CUDA(ctx, cudaMemcpy(nodes.data(), ctx->nodes, ctx->nodeBytes, cudaMemcpyDeviceToHost));
    if (ctx->woopBytes) B2RT_CUDA(ctx, cudaMemcpy(woop.data(), ctx->woop, ctx->woopBytes, cudaMemcpyDeviceToHost));
    if (ctx->triIndexBytes) B2RT_CUDA(ctx, cudaMemcpy(tri.data(), ctx->triIndex, ctx->triIndexBytes, cudaMemcpyDeviceToHost));
    Wide8 w;
    const char* err = "";
    if (!buildWide8(nodes.data(), nodes.size() / 16, woop.data(), tri.data(), woop.size() / 4, w, &err, greedy))
        return setError(ctx, B2RT_ERR_ARG, "wide BVH build failed: %s", err);
    return uploadWide(ctx, w.nodes.data(), w.nodes.size() * 4, w.woop.data(), w.woop.size() * 4, w.triIndex.data(),
                      w.triIndex.size() * 4, w.numNodes, w.numTris, w.maxDepth, greedy, false);
}

// The three arrays of the wide form go into ONE allocation (256-byte aligned parts), so that a single persisting-L2
// access window can cover them (launchTraceWide).  srcIsDevice: the arrays are device memory already (b2rt_mg_broadcast_bvh).
int uploadWide(b2rt_ctx* ctx, const void* nodes, size_t nodeBytes, const void* woop, size_t woopBytes, const void* tri,
               size_t triBytes, int numNodes, int numTris, int maxDepth, bool greedy, bool srcIsDevice)
{
    freeWide(ctx);
    auto up = [](size_t v) { return (v + 255) & ~size_t(255); };
    const size_t total = up(nodeBytes) + up(woopBytes) + up(triBytes + 16);
    char* base = nullptr;
    B2RT_CUDA(ctx, cudaMalloc(&base, total));
    ctx->wideNodes = reinterpret_cast<float4*>(base);
    ctx->wideWoop = reinterpret_cast<float4*>(base + up(nodeBytes));
    ctx->wideTriIndex = reinterpret_cast<int32_t*>(base + up(nodeBytes) + up(woopBytes));
    ctx->wideBytes = total;
    const cudaMemcpyKind kind = srcIsDevice ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    B2RT_CUDA(ctx, cudaMemcpy(ctx->wideNodes, nodes, nodeBytes, kind));
    B2RT_CUDA(ctx, cudaMemcpy(ctx->wideWoop, woop, woopBytes, kind));
    if (triBytes) B2RT_CUDA(ctx, cudaMemcpy(ctx->wideTriIndex, tri, triBytes, kind));
    ctx->wideNumNodes = numNodes; ctx->wideNumTris = numTris; ctx->wideMaxDepth = maxDepth;
    ctx->wideGreedy = greedy;
    ctx->hasWide = true;
    return B2RT_OK;
}

}  // namespace b2rt

using namespace b2rt;

extern "C" {

int b2rt_create(int device, b2rt_ctx** out)
{
    if (!out) return setError(nullptr, B2RT_ERR_ARG, "b2rt_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return setError(nullptr, B2RT_ERR_NO_DEVICE, "b2rt_create: no CUDA device (%s)", cudaGetErrorString(e));
    if (device < 0 || device >= count)
        return setError(nullptr, B2RT_ERR_ARG, "b2rt_create: device %d out of range [0,%d)", device, count);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess)
        return setError(nullptr, B2RT_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    if (prop.major != 10)
        return setError(nullptr, B2RT_ERR_NO_DEVICE,
                        "b2rt_create: device %d is sm_%d%d; this library carries sm_100a code only (no fallback)",
                        device, prop.major, prop.minor);
    b2rt_ctx* ctx = new b2rt_ctx();
    ctx->device = device;
    ctx->smCount = prop.multiProcessorCount;
    ctx->l2Bytes = (size_t)prop.l2CacheSize;
    ctx->ccMajor = prop.major; ctx->ccMinor = prop.minor;
    if ((e = cudaSetDevice(device)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->warpCounter, b2rt_ctx::kLaunchSlots * 128)) != cudaSuccess ||
        (e = cudaMalloc(&ctx->statsDev, (16 + 3 * kTimelineWarps) * sizeof(unsigned long long))) != cudaSuccess ||
        (e = cudaMalloc(&ctx->aabbDev, 8 * sizeof(float))) != cudaSuccess ||
        (e = cudaMalloc(&ctx->countDev, 2 * sizeof(long long))) != cudaSuccess ||
        (e = cudaMemset(ctx->countDev, 0, 2 * sizeof(long long))) != cudaSuccess ||
        (e = cudaHostAlloc(&ctx->countHost, sizeof(long long), cudaHostAllocMapped | cudaHostAllocPortable)) != cudaSuccess ||
        (e = cudaEventCreate(&ctx->ev0)) != cudaSuccess || (e = cudaEventCreate(&ctx->ev1)) != cudaSuccess) {
        setError(nullptr, B2RT_ERR_CUDA, "b2rt_create: %s", cudaGetErrorString(e));
        b2rt_destroy(ctx);
        return B2RT_ERR_CUDA;
    }
    cudaMemset(ctx->statsDev, 0, (16 + 3 * kTimelineWarps) * sizeof(unsigned long long));
    *out = ctx;
    return B2RT_OK;
}

void b2rt_destroy(b2rt_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaFree(ctx->nodes); cudaFree(ctx->woop); cudaFree(ctx->triIndex);
    cudaFree(ctx->wideNodes);
    cudaFree(ctx->shareList); cudaFree(ctx->warpCounter); cudaFree(ctx->statsDev); cudaFree(ctx->scratch);
    for (auto& slot : ctx->hostSlot) cudaFree(slot.stage);
    cudaFree(ctx->aabbDev); cudaFree(ctx->countDev); cudaFreeHost(ctx->countHost);
    if (ctx->sIn) {
        cudaStreamDestroy(ctx->sIn); cudaStreamDestroy(ctx->sRun); cudaStreamDestroy(ctx->sOut);
        for (auto& slot : ctx->hostSlot) {
            for (int i = 0; i < b2rt_ctx::kHostChunks; i++) { cudaEventDestroy(slot.evIn[i]); cudaEventDestroy(slot.evK0[i]); cudaEventDestroy(slot.evK1[i]); }
            cudaEventDestroy(slot.evT0); cudaEventDestroy(slot.evT1);
        }
    }
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    delete ctx;
}

const char* b2rt_last_error(const b2rt_ctx* ctx) { return ctx ? ctx->lastError.c_str() : g_createError.c_str(); }
int64_t b2rt_launch_count(const b2rt_ctx* ctx) { return ctx ? ctx->launches : 0; }

int b2rt_set_trace_config(b2rt_ctx* ctx, const b2rt_trace_config* cfg)
{
    if (!ctx || !cfg) return setError(ctx, B2RT_ERR_ARG, "b2rt_set_trace_config: NULL argument");
    ctx->cfg = *cfg;
    return B2RT_OK;
}

int b2rt_device_info(const b2rt_ctx* ctx, int* sm, int* major, int* minor)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (sm) *sm = ctx->smCount;
    if (major) *major = ctx->ccMajor;
    if (minor) *minor = ctx->ccMinor;
    return B2RT_OK;
}

int b2rt_set_bvh(b2rt_ctx* ctx, const void* nodes, size_t nodeBytes, const void* woop, size_t woopBytes,
                 const int32_t* triIndex, size_t triIndexBytes, int layout, int srcIsDevice)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (layout != B2RT_LAYOUT_COMPACT2) return setError(ctx, B2RT_ERR_LAYOUT, "CudaTracer: Incorrect BVH layout!");
    if (!nodes || !woop || !triIndex || nodeBytes < 64 || nodeBytes % 64 || woopBytes % 16 ||
        triIndexBytes * 4 != woopBytes)
        return setError(ctx, B2RT_ERR_ARG, "b2rt_set_bvh: inconsistent array sizes (nodes %zu, woop %zu, triIndex %zu)",
                        nodeBytes, woopBytes, triIndexBytes);
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaFree(ctx->nodes); cudaFree(ctx->woop); cudaFree(ctx->triIndex);
    ctx->nodes = nullptr; ctx->woop = nullptr; ctx->triIndex = nullptr; ctx->hasBvh = false;
    freeWide(ctx);

    // nodes: host round trip for the re-layout (one-off, at most tens of MB)
    std::vector<float> in(nodeBytes / 4), out;
    B2RT_CUDA(ctx, cudaMemcpy(in.data(), nodes, nodeBytes, srcIsDevice ? cudaMemcpyDeviceToHost : cudaMemcpyHostToHost));
    int topCount = 0;
    relayoutNodes(in, out, topCount);
    B2RT_CUDA(ctx, cudaMalloc(&ctx->nodes, nodeBytes));
    B2RT_CUDA(ctx, cudaMemcpy(ctx->nodes, out.data(), nodeBytes, cudaMemcpyHostToDevice));
    // Woop rows: the leaf loop fetches three rows before it looks at the terminator, so the copy
    // carries padding rows past the last terminator (the reference relies on texture fetches
    // returning zero out of range, kepler_dynamic_fetch.cu:328-334).  The first padding row is
    // itself a terminator: row woopBytes/16 is a leaf without triangles, which the any-hit kernel leaves
    // in place of stack entries it hands to other lanes (trace.cu, SHARE).
    const cudaMemcpyKind kind = srcIsDevice ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    B2RT_CUDA(ctx, cudaMalloc(&ctx->woop, woopBytes + kWoopPadBytes));
    B2RT_CUDA(ctx, cudaMemset(reinterpret_cast<char*>(ctx->woop) + woopBytes, 0, kWoopPadBytes));
    {
        const unsigned terminator = 0x80000000u;
        B2RT_CUDA(ctx, cudaMemcpy(reinterpret_cast<char*>(ctx->woop) + woopBytes, &terminator, 4, cudaMemcpyHostToDevice));
    }
    if (woopBytes) B2RT_CUDA(ctx, cudaMemcpy(ctx->woop, woop, woopBytes, kind));
    B2RT_CUDA(ctx, cudaMalloc(&ctx->triIndex, triIndexBytes + 16));
    if (triIndexBytes) B2RT_CUDA(ctx, cudaMemcpy(ctx->triIndex, triIndex, triIndexBytes, kind));
    ctx->nodeBytes = nodeBytes; ctx->woopBytes = woopBytes; ctx->triIndexBytes = triIndexBytes;
    ctx->topNodesAvail = topCount;
    ctx->hasBvh = true;
    return B2RT_OK;
}

int b2rt_get_bvh(const b2rt_ctx* ctx, const void** nodes, size_t* nodeBytes, const void** woop, size_t* woopBytes,
                 const int32_t** triIndex, size_t* triIndexBytes)
{
    if (!ctx || !ctx->hasBvh) return B2RT_ERR_NO_BVH;
    if (nodes) *nodes = ctx->nodes;
    if (nodeBytes) *nodeBytes = ctx->nodeBytes;
    if (woop) *woop = ctx->woop;
    if (woopBytes) *woopBytes = ctx->woopBytes;
    if (triIndex) *triIndex = ctx->triIndex;
    if (triIndexBytes) *triIndexBytes = ctx->triIndexBytes;
    return B2RT_OK;
}

int b2rt_trace(b2rt_ctx* ctx, const void* rays, void* results, int64_t numRays, unsigned flags, void* stream, float* ms)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (ms) *ms = 0.0f;
    if (numRays == 0) return B2RT_OK;                                   // CudaTracer.cc:123-125
    if (!ctx->hasBvh) return setError(ctx, B2RT_ERR_NO_BVH, "CudaTracer: No BVH!");
    if (numRays < 0 || numRays > 0x7fffffff || !rays || !results)
        return setError(ctx, B2RT_ERR_ARG, "b2rt_trace: bad ray buffer (n=%lld)", (long long)numRays);
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (ms) B2RT_CUDA(ctx, cudaEventRecord(ctx->ev0, s));
    int rc = launchTrace(ctx, rays, results, numRays, flags, s);
    if (rc) return rc;
    if (ms) {
        B2RT_CUDA(ctx, cudaEventRecord(ctx->ev1, s));
        B2RT_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
        B2RT_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    }
    return B2RT_OK;
}

// Wait for one in-flight host batch and add its times to the running totals.
static int finishHostSlot(b2rt_ctx* ctx, b2rt_ctx::HostSlot& slot, float* msKernel, cudaEvent_t* first, cudaEvent_t* last)
{
    if (!slot.inFlight) return B2RT_OK;
    slot.inFlight = false;
    B2RT_CUDA(ctx, cudaEventSynchronize(slot.evT1));
    if (msKernel)
        for (int c = 0; c < slot.chunks; c++) { float t = 0.0f; cudaEventElapsedTime(&t, slot.evK0[c], slot.evK1[c]); *msKernel += t; }
    if (first && !*first) *first = slot.evT0;
    if (last) *last = slot.evT1;
    return B2RT_OK;
}

int b2rt_trace_host_begin(b2rt_ctx* ctx, const void* raysHost, void* resultsHost, int64_t numRays, unsigned flags)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (numRays == 0) return B2RT_OK;
    if (!ctx->hasBvh) return setError(ctx, B2RT_ERR_NO_BVH, "CudaTracer: No BVH!");
    if (numRays < 0 || numRays > 0x7fffffff || !raysHost || !resultsHost)
        return setError(ctx, B2RT_ERR_ARG, "b2rt_trace_host: bad ray buffer (n=%lld)", (long long)numRays);
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    // Three streams, up to kHostChunks chunks per batch: the upload of chunk i+1, the trace of chunk i and the download of
    // chunk i-1 overlap (PCIe is full duplex), so a batch costs about max(upload, download) instead of their sum + the
    // kernel; with a second batch begun before the first is waited for, its upload also hides the first one's last
    // kernel and download.
    if (!ctx->sIn) {
        B2RT_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->sIn, cudaStreamNonBlocking));
        B2RT_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->sRun, cudaStreamNonBlocking));
        B2RT_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->sOut, cudaStreamNonBlocking));
        for (auto& slot : ctx->hostSlot) {
            for (int i = 0; i < b2rt_ctx::kHostChunks; i++) {
                B2RT_CUDA(ctx, cudaEventCreateWithFlags(&slot.evIn[i], cudaEventDisableTiming));
                B2RT_CUDA(ctx, cudaEventCreate(&slot.evK0[i]));
                B2RT_CUDA(ctx, cudaEventCreate(&slot.evK1[i]));
            }
            B2RT_CUDA(ctx, cudaEventCreate(&slot.evT0));
            B2RT_CUDA(ctx, cudaEventCreate(&slot.evT1));
        }
    }
    b2rt_ctx::HostSlot& slot = ctx->hostSlot[ctx->nextHostSlot];
    ctx->nextHostSlot = (ctx->nextHostSlot + 1) % b2rt_ctx::kHostSlots;
    int rc = finishHostSlot(ctx, slot, nullptr, nullptr, nullptr);       // both slots busy: the older batch has to drain first
    if (rc) return rc;
    const size_t rayBytes = (size_t)numRays * sizeof(Ray), resBytes = (size_t)numRays * sizeof(RayResult);
    rc = ensureStage(ctx, slot, rayBytes + resBytes);
    if (rc) return rc;
    // 256 Ki rays per chunk: measured best on C2 (64 Ki 876, 128 Ki 1 055, 192 Ki 1 225, 256 Ki 1 275, 400 K 1 138 Mrays/s
    // end to end) -- smaller chunks make the launches tail-bound and slower than their own upload.  At most kHostChunks (32)
    // chunks per batch: what a batch costs beyond its upload is the LAST chunk's trace and download, so a 64 Mi-ray batch goes
    // in 32 chunks of 2 Mi rays rather than 8 of 8 Mi (see DESIGN.md section 5 for the measured difference).
    int chunks = (int)std::min<int64_t>(b2rt_ctx::kHostChunks, std::max<int64_t>(1, numRays / (256 * 1024)));
    const int64_t per = ((numRays + chunks - 1) / chunks + 31) & ~int64_t(31);
    chunks = (int)((numRays + per - 1) / per);
    char* dRays = static_cast<char*>(slot.stage);
    char* dRes = dRays + rayBytes;
    const char* hRays = static_cast<const char*>(raysHost);
    char* hRes = static_cast<char*>(resultsHost);
    const bool clearPads = !(flags & (B2RT_TRACE_WRITE_BARY | B2RT_TRACE_STATS | B2RT_TRACE_WATERTIGHT));
    cudaError_t e = cudaEventRecord(slot.evT0, ctx->sIn);
    for (int c = 0; c < chunks && e == cudaSuccess && rc == 0; c++) {
        const int64_t lo = c * per, n = std::min<int64_t>(per, numRays - lo);
        e = cudaMemcpyAsync(dRays + lo * sizeof(Ray), hRays + lo * sizeof(Ray), (size_t)n * sizeof(Ray), cudaMemcpyHostToDevice, ctx->sIn);
        if (e == cudaSuccess) e = cudaEventRecord(slot.evIn[c], ctx->sIn);
        // the kernel writes only id and t without WRITE_BARY (as CudaTracerKernels.hh:197): the pad words that
        // travel back to the host are zero, not stale staging memory
        if (e == cudaSuccess && clearPads) e = cudaMemsetAsync(dRes + lo * sizeof(RayResult), 0, (size_t)n * sizeof(RayResult), ctx->sRun);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->sRun, slot.evIn[c], 0);
        if (e == cudaSuccess) e = cudaEventRecord(slot.evK0[c], ctx->sRun);
        if (e == cudaSuccess) rc = launchTrace(ctx, dRays + lo * sizeof(Ray), dRes + lo * sizeof(RayResult), n, flags, ctx->sRun);
        if (e == cudaSuccess && rc == 0) e = cudaEventRecord(slot.evK1[c], ctx->sRun);
        if (e == cudaSuccess && rc == 0) e = cudaStreamWaitEvent(ctx->sOut, slot.evK1[c], 0);
        if (e == cudaSuccess && rc == 0)
            e = cudaMemcpyAsync(hRes + lo * sizeof(RayResult), dRes + lo * sizeof(RayResult), (size_t)n * sizeof(RayResult), cudaMemcpyDeviceToHost, ctx->sOut);
    }
    if (e == cudaSuccess && rc == 0) e = cudaEventRecord(slot.evT1, ctx->sOut);
    if (rc || e != cudaSuccess) {
        cudaStreamSynchronize(ctx->sOut); cudaStreamSynchronize(ctx->sRun); cudaStreamSynchronize(ctx->sIn);
        if (rc) return rc;
        return setError(ctx, B2RT_ERR_CUDA, "b2rt_trace_host: %s", cudaGetErrorString(e));
    }
    slot.chunks = chunks;
    slot.inFlight = true;
    return B2RT_OK;
}

int b2rt_trace_host_end(b2rt_ctx* ctx, float* msKernel, float* msTotal)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (msKernel) *msKernel = 0.0f;
    if (msTotal) *msTotal = 0.0f;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaEvent_t first = nullptr, last = nullptr;
    // oldest first: nextHostSlot is the slot the next begin would take, i.e. the older of the two
    for (int k = 0; k < b2rt_ctx::kHostSlots; k++) {
        const int rc = finishHostSlot(ctx, ctx->hostSlot[(ctx->nextHostSlot + k) % b2rt_ctx::kHostSlots], msKernel, &first, &last);
        if (rc) return rc;
    }
    if (msTotal && first && last) cudaEventElapsedTime(msTotal, first, last);
    return B2RT_OK;
}

int b2rt_trace_host(b2rt_ctx* ctx, const void* raysHost, void* resultsHost, int64_t numRays, unsigned flags,
                    float* msKernel, float* msTotal)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (msKernel) *msKernel = 0.0f;
    if (msTotal) *msTotal = 0.0f;
    int rc = b2rt_trace_host_end(ctx, nullptr, nullptr);         // batches begun earlier are not part of this call's times
    if (rc) return rc;
    rc = b2rt_trace_host_begin(ctx, raysHost, resultsHost, numRays, flags);
    if (rc) return rc;
    return b2rt_trace_host_end(ctx, msKernel, msTotal);
}

int b2rt_count_work(b2rt_ctx* ctx, const void* rays, int64_t numRays, int anyHit, b2rt_work_counters* out, void* stream)
{
    if (!ctx || !out) return B2RT_ERR_ARG;
    memset(out, 0, sizeof(*out));
    if (numRays == 0) return B2RT_OK;
    if (!ctx->hasBvh) return setError(ctx, B2RT_ERR_NO_BVH, "CudaTracer: No BVH!");
    if (numRays < 0 || numRays > 0x7fffffff || !rays) return setError(ctx, B2RT_ERR_ARG, "b2rt_count_work: bad ray buffer");
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    return countWork(ctx, rays, numRays, anyHit, out, static_cast<cudaStream_t>(stream));
}

int b2rt_get_trace_stats(b2rt_ctx* ctx, b2rt_trace_stats* out)
{
    if (!ctx || !out) return B2RT_ERR_ARG;
    unsigned long long h[16];
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    B2RT_CUDA(ctx, cudaDeviceSynchronize());
    B2RT_CUDA(ctx, cudaMemcpy(h, ctx->statsDev, sizeof(h), cudaMemcpyDeviceToHost));
    out->node_iters = h[0]; out->node_lanes = h[1]; out->tri_iters = h[2]; out->tri_lanes = h[3];
    out->fetches = h[4]; out->rays = h[5];
    out->warps = h[12];
    out->ns_pool_empty = out->ns_mean_exit = out->ns_last_exit = 0;
    out->ns_exit_p50 = out->ns_exit_p90 = out->ns_exit_p99 = 0;
    out->ns_straggler_fetch = out->straggler_node_iters = out->straggler_tri_iters = 0;
    out->shares = h[13];
    if (h[12] && h[8] != ~0ull) {
        const unsigned long long start40 = h[8] & 0xffffffffffull;
        out->ns_last_exit = h[11] - h[8];
        out->ns_pool_empty = (h[9] != ~0ull) ? h[9] - h[8] : 0;
        out->ns_mean_exit = h[10] / h[12] - start40;        // valid unless the 40-bit timer wrapped during the launch (18 min period)
        const size_t n = h[12] < (unsigned long long)kTimelineWarps ? (size_t)h[12] : (size_t)kTimelineWarps;
        std::vector<unsigned long long> exits(n);
        B2RT_CUDA(ctx, cudaMemcpy(exits.data(), ctx->statsDev + 16, n * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
        {
            const size_t last = std::max_element(exits.begin(), exits.end()) - exits.begin();
            unsigned long long fetch = 0, iters = 0;
            B2RT_CUDA(ctx, cudaMemcpy(&fetch, ctx->statsDev + 16 + kTimelineWarps + last, 8, cudaMemcpyDeviceToHost));
            B2RT_CUDA(ctx, cudaMemcpy(&iters, ctx->statsDev + 16 + 2 * kTimelineWarps + last, 8, cudaMemcpyDeviceToHost));
            out->ns_straggler_fetch = fetch > h[8] ? fetch - h[8] : 0;
            out->straggler_node_iters = iters & 0xffffffffull;
            out->straggler_tri_iters = iters >> 32;
        }
        std::sort(exits.begin(), exits.end());
        auto pct = [&](double q) { const unsigned long long v = exits[(size_t)(q * (n - 1))]; return v > h[8] ? v - h[8] : 0ull; };
        out->ns_exit_p50 = pct(0.50); out->ns_exit_p90 = pct(0.90); out->ns_exit_p99 = pct(0.99);
    }
    return B2RT_OK;
}

}  // extern "C"
