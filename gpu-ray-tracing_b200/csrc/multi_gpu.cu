// multi_gpu.cu -- b2rt_mg_*: BVH replicated on the GPUs of one box, ray batches sharded (include/b2rt_mg.h).
// No counterpart in the reference (single-GPU); contract = BASELINE.json north_star + SURVEY.md 8(b)/(e).
// One host thread drives every device: kernel launches and copies are asynchronous on one stream per device, NCCL calls
// are issued inside ncclGroupStart / ncclGroupEnd.  NCCL is bound with dlopen so that libb2rt.so does not depend on it.
#include "common.cuh"
#include "../../include/b2rt_mg.h"

#include <dlfcn.h>
#include <stdarg.h>
#include <nccl.h>

#include <algorithm>
#include <vector>

namespace b2rt {

constexpr int kMaxMgDevices = 16;

// ---- NCCL, bound at run time ------------------------------------------------------------------
struct Nccl {
    void* lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
};
static Nccl g_nccl;
static std::string g_mgCreateError;

static bool loadNccl(std::string& why)
{
    if (g_nccl.lib) return true;
    // a copy already in the process (torch's) is found first: same SONAME
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { why = std::string("cannot load libnccl.so.2: ") + dlerror(); return false; }
    Nccl n;
    n.lib = h;
#define B2RT_NCCL_SYM(field, name)                                                        \
    *reinterpret_cast<void**>(&n.field) = dlsym(h, name);                                 \
    if (!n.field) { why = std::string("libnccl lacks ") + name; dlclose(h); return false; }
    B2RT_NCCL_SYM(CommInitAll, "ncclCommInitAll")
    B2RT_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    B2RT_NCCL_SYM(GroupStart, "ncclGroupStart")
    B2RT_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    B2RT_NCCL_SYM(Broadcast, "ncclBroadcast")
    B2RT_NCCL_SYM(AllReduce, "ncclAllReduce")
    B2RT_NCCL_SYM(GetErrorString, "ncclGetErrorString")
    B2RT_NCCL_SYM(GetVersion, "ncclGetVersion")
#undef B2RT_NCCL_SYM
    g_nccl = n;
    return true;
}

}  // namespace b2rt

struct b2rt_mg {
    int n = 0;
    int device[b2rt::kMaxMgDevices] = {};
    b2rt_ctx* ctx[b2rt::kMaxMgDevices] = {};
    cudaStream_t stream[b2rt::kMaxMgDevices] = {};
    ncclComm_t comm[b2rt::kMaxMgDevices] = {};
    cudaEvent_t ev0[b2rt::kMaxMgDevices] = {}, ev1[b2rt::kMaxMgDevices] = {}, ev2[b2rt::kMaxMgDevices] = {};
    long long* hitsDev[b2rt::kMaxMgDevices] = {};           // per device: [0] own count, [1] sum over devices
    void* stage[b2rt::kMaxMgDevices] = {};                  // b2rt_mg_trace_host: rays, then results of the device's slice
    size_t stageBytes[b2rt::kMaxMgDevices] = {};
    bool haveComm = false;
    std::string lastError;
};

namespace b2rt {

static int mgError(b2rt_mg* mg, int code, const char* fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (mg) mg->lastError = buf; else g_mgCreateError = buf;
    return code;
}

#define B2RT_MG_CUDA(mg, call)                                                                          \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess)                                                                          \
            return b2rt::mgError((mg), B2RT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)
#define B2RT_MG_NCCL(mg, call)                                                                          \
    do {                                                                                                \
        ncclResult_t r_ = (call);                                                                       \
        if (r_ != ncclSuccess)                                                                          \
            return b2rt::mgError((mg), B2RT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, b2rt::g_nccl.GetErrorString(r_), __FILE__, __LINE__); \
    } while (0)

// id >= 0 count of one shard into a device word (no host round trip: the count feeds the all-reduce)
__global__ void __launch_bounds__(256) mgCountHitsKernel(const int4* __restrict__ results, long long n, long long* out)
{
    unsigned c = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        c += (__ldg(reinterpret_cast<const int*>(results + i)) >= 0) ? 1u : 0u;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(reinterpret_cast<unsigned long long*>(out), (unsigned long long)c);
}

}  // namespace b2rt

using namespace b2rt;

extern "C" {

const char* b2rt_mg_last_error(const b2rt_mg* mg) { return mg ? mg->lastError.c_str() : g_mgCreateError.c_str(); }
int b2rt_mg_num_devices(const b2rt_mg* mg) { return mg ? mg->n : 0; }
b2rt_ctx* b2rt_mg_context(b2rt_mg* mg, int rank) { return (mg && rank >= 0 && rank < mg->n) ? mg->ctx[rank] : nullptr; }
void* b2rt_mg_stream(b2rt_mg* mg, int rank) { return (mg && rank >= 0 && rank < mg->n) ? mg->stream[rank] : nullptr; }

void b2rt_mg_destroy(b2rt_mg* mg)
{
    if (!mg) return;
    for (int r = 0; r < mg->n; r++) {
        if (!mg->ctx[r]) continue;
        cudaSetDevice(mg->device[r]);
        cudaDeviceSynchronize();
        if (mg->haveComm && mg->comm[r]) g_nccl.CommDestroy(mg->comm[r]);
        if (mg->ev0[r]) cudaEventDestroy(mg->ev0[r]);
        if (mg->ev1[r]) cudaEventDestroy(mg->ev1[r]);
        if (mg->ev2[r]) cudaEventDestroy(mg->ev2[r]);
        cudaFree(mg->hitsDev[r]);
        cudaFree(mg->stage[r]);
        if (mg->stream[r]) cudaStreamDestroy(mg->stream[r]);
        b2rt_destroy(mg->ctx[r]);
    }
    delete mg;
}

int b2rt_mg_create(int num_devices, const int* devices, b2rt_mg** out)
{
    if (!out) return mgError(nullptr, B2RT_ERR_ARG, "b2rt_mg_create: out is NULL");
    *out = nullptr;
    if (num_devices < 1 || num_devices > kMaxMgDevices)
        return mgError(nullptr, B2RT_ERR_ARG, "b2rt_mg_create: %d devices (1..%d supported)", num_devices, kMaxMgDevices);
    std::string why;
    if (!loadNccl(why)) return mgError(nullptr, B2RT_ERR_NO_DEVICE, "b2rt_mg_create: %s", why.c_str());
    b2rt_mg* mg = new b2rt_mg();
    mg->n = num_devices;
    for (int r = 0; r < num_devices; r++) mg->device[r] = devices ? devices[r] : r;
    for (int r = 0; r < num_devices; r++)
        for (int q = 0; q < r; q++)
            if (mg->device[q] == mg->device[r]) {
                const int twice = mg->device[r];
                delete mg;
                return mgError(nullptr, B2RT_ERR_ARG, "b2rt_mg_create: device %d listed twice", twice);
            }
    auto fail = [&](int code, const std::string& msg) { b2rt_mg_destroy(mg); return mgError(nullptr, code, "b2rt_mg_create: %s", msg.c_str()); };
    for (int r = 0; r < num_devices; r++) {
        const int rc = b2rt_create(mg->device[r], &mg->ctx[r]);
        if (rc) { const std::string m = b2rt_last_error(nullptr); mg->n = r; return fail(rc, m); }
        cudaError_t e = cudaSetDevice(mg->device[r]);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&mg->stream[r], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreate(&mg->ev0[r]);
        if (e == cudaSuccess) e = cudaEventCreate(&mg->ev1[r]);
        if (e == cudaSuccess) e = cudaEventCreate(&mg->ev2[r]);
        if (e == cudaSuccess) e = cudaMalloc(&mg->hitsDev[r], 2 * sizeof(long long));
        if (e != cudaSuccess) { mg->n = r + 1; return fail(B2RT_ERR_CUDA, cudaGetErrorString(e)); }
    }
    const ncclResult_t nr = g_nccl.CommInitAll(mg->comm, num_devices, mg->device);
    if (nr != ncclSuccess) return fail(B2RT_ERR_CUDA, std::string("ncclCommInitAll: ") + g_nccl.GetErrorString(nr));
    mg->haveComm = true;
    *out = mg;
    return B2RT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// scene: built once on the first device, broadcast into the other contexts' private arrays
// ---------------------------------------------------------------------------------------------
extern "C" int b2rt_mg_broadcast_bvh(b2rt_mg* mg, const void* nodes, size_t nodeBytes, const void* woop, size_t woopBytes,
                                     const int32_t* triIndex, size_t triIndexBytes, int layout, int srcIsDevice)
{
    if (!mg) return B2RT_ERR_ARG;
    b2rt_ctx* root = mg->ctx[0];
    // the first device: the caller's arrays -> private re-laid copy (api.cu:b2rt_set_bvh), and the 8-wide form if the default
    // policy would trace with it (trace.cu:launchTrace: nodes + Woop rows above 48 MB)
    int rc = b2rt_set_bvh(root, nodes, nodeBytes, woop, woopBytes, triIndex, triIndexBytes, layout, srcIsDevice);
    if (rc) return mgError(mg, rc, "b2rt_mg_broadcast_bvh: %s", b2rt_last_error(root));
    const bool wide = root->nodeBytes + root->woopBytes > (size_t(48) << 20);
    if (wide) {
        rc = ensureWide(root);
        if (rc) return mgError(mg, rc, "b2rt_mg_broadcast_bvh: %s", b2rt_last_error(root));
    }
    if (mg->n == 1) return B2RT_OK;
    const size_t woopAll = root->woopBytes + kWoopPadBytes, triAll = root->triIndexBytes + 16;
    for (int r = 1; r < mg->n; r++) {
        b2rt_ctx* c = mg->ctx[r];
        B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[r]));
        cudaFree(c->nodes); cudaFree(c->woop); cudaFree(c->triIndex);
        c->nodes = nullptr; c->woop = nullptr; c->triIndex = nullptr; c->hasBvh = false;
        freeWide(c);
        B2RT_MG_CUDA(mg, cudaMalloc(&c->nodes, root->nodeBytes));
        B2RT_MG_CUDA(mg, cudaMalloc(&c->woop, woopAll));
        B2RT_MG_CUDA(mg, cudaMalloc(&c->triIndex, triAll));
        if (wide) {
            char* base = nullptr;
            B2RT_MG_CUDA(mg, cudaMalloc(&base, root->wideBytes));
            c->wideNodes = reinterpret_cast<float4*>(base);
            c->wideWoop = reinterpret_cast<float4*>(base + (reinterpret_cast<char*>(root->wideWoop) - reinterpret_cast<char*>(root->wideNodes)));
            c->wideTriIndex = reinterpret_cast<int32_t*>(base + (reinterpret_cast<char*>(root->wideTriIndex) - reinterpret_cast<char*>(root->wideNodes)));
            c->wideBytes = root->wideBytes;
        }
    }
    // ncclBroadcast x3 (+1 for the wide form, one allocation), root = rank 0, in place on the root
    struct Part { void* ptr[kMaxMgDevices]; size_t bytes; };
    std::vector<Part> parts;
    auto add = [&](size_t bytes, auto field) {
        Part p; p.bytes = bytes;
        for (int r = 0; r < mg->n; r++) p.ptr[r] = field(mg->ctx[r]);
        if (bytes) parts.push_back(p);
    };
    add(root->nodeBytes, [](b2rt_ctx* c) -> void* { return c->nodes; });
    add(woopAll, [](b2rt_ctx* c) -> void* { return c->woop; });
    add(triAll, [](b2rt_ctx* c) -> void* { return c->triIndex; });
    if (wide) add(root->wideBytes, [](b2rt_ctx* c) -> void* { return c->wideNodes; });
    B2RT_MG_NCCL(mg, g_nccl.GroupStart());
    for (const Part& p : parts)
        for (int r = 0; r < mg->n; r++)
            B2RT_MG_NCCL(mg, g_nccl.Broadcast(p.ptr[r], p.ptr[r], p.bytes, ncclChar, 0, mg->comm[r], mg->stream[r]));
    B2RT_MG_NCCL(mg, g_nccl.GroupEnd());
    for (int r = 0; r < mg->n; r++) {
        B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[r]));
        B2RT_MG_CUDA(mg, cudaStreamSynchronize(mg->stream[r]));
    }
    for (int r = 1; r < mg->n; r++) {
        b2rt_ctx* c = mg->ctx[r];
        c->nodeBytes = root->nodeBytes; c->woopBytes = root->woopBytes; c->triIndexBytes = root->triIndexBytes;
        c->topNodesAvail = root->topNodesAvail;
        c->hasBvh = true;
        if (wide) {
            c->wideNumNodes = root->wideNumNodes; c->wideNumTris = root->wideNumTris; c->wideMaxDepth = root->wideMaxDepth;
            c->wideGreedy = root->wideGreedy; c->hasWide = true;
        }
    }
    return B2RT_OK;
}

// ---------------------------------------------------------------------------------------------
// trace: every device its shard; gather of the hit records and sum of the hit counts on request
// ---------------------------------------------------------------------------------------------
static int traceShards(b2rt_mg* mg, const void* const* rays, void* const* results, const int64_t* numRays, unsigned flags,
                       void* const* gathered, b2rt_mg_stats* stats)
{
    int64_t total = 0;
    for (int r = 0; r < mg->n; r++) {
        if (numRays[r] < 0 || numRays[r] > 0x7fffffff || (numRays[r] && (!rays[r] || !results[r])))
            return mgError(mg, B2RT_ERR_ARG, "b2rt_mg_trace: bad shard %d (n=%lld)", r, (long long)numRays[r]);
        if (gathered && !gathered[r]) return mgError(mg, B2RT_ERR_ARG, "b2rt_mg_trace: gathered_dev[%d] is NULL", r);
        total += numRays[r];
    }
    flags &= ~B2RT_TRACE_STATS;                      // the instrumented kernel's counters are per context: not summed here
    for (int r = 0; r < mg->n; r++) {
        B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[r]));
        B2RT_MG_CUDA(mg, cudaEventRecord(mg->ev0[r], mg->stream[r]));
        if (numRays[r]) {
            if (!mg->ctx[r]->hasBvh) return mgError(mg, B2RT_ERR_NO_BVH, "CudaTracer: No BVH!");
            const int rc = launchTrace(mg->ctx[r], rays[r], results[r], numRays[r], flags, mg->stream[r]);
            if (rc) return mgError(mg, rc, "b2rt_mg_trace (device %d): %s", mg->device[r], b2rt_last_error(mg->ctx[r]));
        }
        B2RT_MG_CUDA(mg, cudaEventRecord(mg->ev1[r], mg->stream[r]));
        // hit count of the shard, on the device (feeds the all-reduce)
        B2RT_MG_CUDA(mg, cudaMemsetAsync(mg->hitsDev[r], 0, 2 * sizeof(long long), mg->stream[r]));
        if (numRays[r]) {
            const unsigned grid = (unsigned)std::min<long long>((numRays[r] + 255) / 256, (long long)mg->ctx[r]->smCount * 8);
            mgCountHitsKernel<<<grid, 256, 0, mg->stream[r]>>>(static_cast<const int4*>(results[r]), numRays[r], mg->hitsDev[r]);
            B2RT_MG_CUDA(mg, cudaGetLastError());
            mg->ctx[r]->launches++;
        }
    }
    // gather (one broadcast per non-empty shard, root = its owner: an all-gather of unequal parts) + sum of the hit counts
    B2RT_MG_NCCL(mg, g_nccl.GroupStart());
    if (gathered) {
        int64_t offset = 0;
        for (int src = 0; src < mg->n; src++) {
            const size_t bytes = (size_t)numRays[src] * sizeof(RayResult);
            if (bytes)
                for (int r = 0; r < mg->n; r++)
                    B2RT_MG_NCCL(mg, g_nccl.Broadcast(results[src], static_cast<char*>(gathered[r]) + offset * sizeof(RayResult), bytes, ncclChar,
                                                      src, mg->comm[r], mg->stream[r]));
            offset += numRays[src];
        }
    }
    for (int r = 0; r < mg->n; r++)
        B2RT_MG_NCCL(mg, g_nccl.AllReduce(mg->hitsDev[r], mg->hitsDev[r] + 1, 1, ncclInt64, ncclSum, mg->comm[r], mg->stream[r]));
    B2RT_MG_NCCL(mg, g_nccl.GroupEnd());
    long long hits = 0;
    for (int r = 0; r < mg->n; r++) {
        B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[r]));
        B2RT_MG_CUDA(mg, cudaEventRecord(mg->ev2[r], mg->stream[r]));
    }
    for (int r = 0; r < mg->n; r++) {
        B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[r]));
        B2RT_MG_CUDA(mg, cudaStreamSynchronize(mg->stream[r]));
    }
    B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[0]));
    B2RT_MG_CUDA(mg, cudaMemcpy(&hits, mg->hitsDev[0] + 1, sizeof(hits), cudaMemcpyDeviceToHost));
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->rays = total; stats->hits = hits;
        stats->ms_trace_min = 3.0e38f;
        for (int r = 0; r < mg->n; r++) {
            float t = 0.0f, g = 0.0f;
            B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[r]));
            B2RT_MG_CUDA(mg, cudaEventElapsedTime(&t, mg->ev0[r], mg->ev1[r]));
            B2RT_MG_CUDA(mg, cudaEventElapsedTime(&g, mg->ev0[r], mg->ev2[r]));
            stats->ms_per_device[r] = t;
            stats->ms_trace_max = std::max(stats->ms_trace_max, t);
            stats->ms_trace_min = std::min(stats->ms_trace_min, t);
            if (gathered) stats->ms_with_gather_max = std::max(stats->ms_with_gather_max, g);
        }
    }
    return B2RT_OK;
}

extern "C" int b2rt_mg_trace(b2rt_mg* mg, const void* const* raysDev, void* const* resultsDev, const int64_t* numRays,
                             unsigned flags, void* const* gatheredDev, b2rt_mg_stats* stats)
{
    if (!mg) return B2RT_ERR_ARG;
    if (!raysDev || !resultsDev || !numRays) return mgError(mg, B2RT_ERR_ARG, "b2rt_mg_trace: NULL argument");
    return traceShards(mg, raysDev, resultsDev, numRays, flags, gatheredDev, stats);
}

extern "C" int b2rt_mg_trace_host(b2rt_mg* mg, const void* raysHost, void* resultsHost, int64_t numRays, unsigned flags,
                                  b2rt_mg_stats* stats)
{
    if (!mg) return B2RT_ERR_ARG;
    if (stats) memset(stats, 0, sizeof(*stats));
    if (numRays == 0) return B2RT_OK;
    if (numRays < 0 || !raysHost || !resultsHost) return mgError(mg, B2RT_ERR_ARG, "b2rt_mg_trace_host: bad ray buffer (n=%lld)", (long long)numRays);
    const int64_t per = (numRays + mg->n - 1) / mg->n;
    const void* rays[kMaxMgDevices]; void* results[kMaxMgDevices]; int64_t count[kMaxMgDevices];
    for (int r = 0; r < mg->n; r++) {
        const int64_t lo = std::min<int64_t>(numRays, r * per), hi = std::min<int64_t>(numRays, lo + per);
        count[r] = hi - lo;
        const size_t need = (size_t)count[r] * (sizeof(Ray) + sizeof(RayResult));
        B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[r]));
        if (need > mg->stageBytes[r]) {
            cudaFree(mg->stage[r]); mg->stage[r] = nullptr; mg->stageBytes[r] = 0;
            B2RT_MG_CUDA(mg, cudaMalloc(&mg->stage[r], need));
            mg->stageBytes[r] = need;
        }
        rays[r] = mg->stage[r];
        results[r] = static_cast<char*>(mg->stage[r]) + (size_t)count[r] * sizeof(Ray);
        if (count[r]) {
            B2RT_MG_CUDA(mg, cudaMemcpyAsync(mg->stage[r], static_cast<const char*>(raysHost) + lo * sizeof(Ray), (size_t)count[r] * sizeof(Ray),
                                             cudaMemcpyHostToDevice, mg->stream[r]));
            // without WRITE_BARY the kernel stores id and t only (CudaTracerKernels.hh:197): the pad words go back as zeros
            B2RT_MG_CUDA(mg, cudaMemsetAsync(results[r], 0, (size_t)count[r] * sizeof(RayResult), mg->stream[r]));
        }
    }
    const int rc = traceShards(mg, rays, results, count, flags, nullptr, stats);
    if (rc) return rc;
    for (int r = 0; r < mg->n; r++) {
        if (!count[r]) continue;
        const int64_t lo = r * per;
        B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[r]));
        B2RT_MG_CUDA(mg, cudaMemcpyAsync(static_cast<char*>(resultsHost) + lo * sizeof(RayResult), results[r], (size_t)count[r] * sizeof(RayResult),
                                         cudaMemcpyDeviceToHost, mg->stream[r]));
    }
    for (int r = 0; r < mg->n; r++) {
        B2RT_MG_CUDA(mg, cudaSetDevice(mg->device[r]));
        B2RT_MG_CUDA(mg, cudaStreamSynchronize(mg->stream[r]));
    }
    return B2RT_OK;
}
