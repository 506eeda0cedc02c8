// static_scratch.cu -- b2rt_count_hits_static: the hit count (misc.cu:countHitsKernel; cuda/RendererKernels.cu:112-162 in the
// reference) for callers that must not allocate -- code under CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc.  No counterpart in the reference.
//
// b2rt_count_hits keeps its accumulator in memory that b2rt_create allocates and hands the result to the host through a
// mapped pinned word, and it needs a context (whose creation allocates).  This variant needs neither: its scratch is the
// file-scope device array below (every GPU that loads the library holds its 33 KB until the process exits), its result goes to
// a device word the CALLER passes, and the call does nothing but one cudaMemsetAsync of 8 bytes and one kernel launch on
// the caller's stream -- no cudaMalloc, no cudaFree, no allocating library routine, no synchronisation.
//
// Scratch layout: [0] = ticket (blocks finished), [32 + b] = hit count of block b.  Block b counts records
// [b * kStaticTile, (b + 1) * kStaticTile); the last block to take a ticket adds the partial counts in block order (so the
// sum does not depend on which block finishes last -- it is an integer sum anyway) and writes the total.  Every partial
// that is read was written by this launch, so only the ticket has to start at zero: it is cleared on the stream before the
// launch and again by the last block.  All calls on a device share the array: two calls that OVERLAP on the GPU (different
// streams, or graph branches that run concurrently) race on it and return garbage; calls on one stream are safe.
#include "common.cuh"

#include <mutex>

namespace b2rt {

constexpr int kStaticTile = 4096;                                          // records per block: 16 per thread
constexpr int kStaticBlocks = B2RT_COUNT_HITS_STATIC_MAX / kStaticTile;    // 4096 partial counts
constexpr int kStaticHeader = 32;                                          // the ticket has a 256-byte line to itself
static_assert(B2RT_COUNT_HITS_STATIC_MAX % kStaticTile == 0, "maximum is a whole number of tiles");

static __device__ __align__(256) unsigned long long g_countScratch[kStaticHeader + kStaticBlocks];

// `scratch` is the DEVICE address of g_countScratch (cudaGetSymbolAddress below).  In host code the array's name is only a
// host-side stand-in for the symbol: passed to a kernel it is not a device address.
__global__ void __launch_bounds__(256) countHitsStaticKernel(const int4* __restrict__ results, long long n,
                                                             unsigned long long* scratch, long long* out)
{
    const long long lo = (long long)blockIdx.x * kStaticTile;
    const long long hi = (lo + kStaticTile < n) ? lo + kStaticTile : n;
    unsigned c = 0;
    for (long long i = lo + threadIdx.x; i < hi; i += 256)
        c += (__ldg(reinterpret_cast<const int*>(results + i)) >= 0) ? 1u : 0u;       // id >= 0 (RendererKernels.cu:139-141)
    c = __reduce_add_sync(0xffffffffu, c);
    __shared__ unsigned s[8];
    __shared__ bool last;
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < 8; w++) t += s[w];
        scratch[kStaticHeader + blockIdx.x] = t;
        __threadfence();
        last = (atomicAdd(&scratch[0], 1ull) == (unsigned long long)gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    // the last block: partial counts of this launch, in block order
    unsigned long long t = 0;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += 256) t += __ldcg(&scratch[kStaticHeader + b]);
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    __shared__ unsigned long long s64[8];
    if ((threadIdx.x & 31) == 0) s64[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long total = 0;
        for (int w = 0; w < 8; w++) total += s64[w];
        *out = (long long)total;
        scratch[0] = 0;                                     // the next call on this stream finds the ticket at zero
    }
}

// device address of the scratch array, looked up once per device with that device current
constexpr int kMaxDevices = 64;
static unsigned long long* g_scratchAddr[kMaxDevices];
static std::mutex g_scratchMutex;

static int scratchForCurrentDevice(unsigned long long** out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return setError(nullptr, B2RT_ERR_NO_DEVICE, "b2rt_count_hits_static: no CUDA device (%s)", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return setError(nullptr, B2RT_ERR_ARG, "b2rt_count_hits_static: device %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_scratchMutex);
    if (!g_scratchAddr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_countScratch);
        if (e != cudaSuccess || !p)
            return setError(nullptr, B2RT_ERR_CUDA, "b2rt_count_hits_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        g_scratchAddr[dev] = static_cast<unsigned long long*>(p);
    }
    *out = g_scratchAddr[dev];
    return B2RT_OK;
}

}  // namespace b2rt

using namespace b2rt;

extern "C" int b2rt_count_hits_static(const void* results_dev, int64_t num_rays, int64_t* hits_dev, void* stream)
{
    if (!hits_dev || num_rays < 0 || (num_rays && !results_dev))
        return setError(nullptr, B2RT_ERR_ARG, "b2rt_count_hits_static: bad arguments (n=%lld)", (long long)num_rays);
    if (num_rays > (int64_t)B2RT_COUNT_HITS_STATIC_MAX)
        return setError(nullptr, B2RT_ERR_ARG, "b2rt_count_hits_static: %lld records exceed the static scratch's maximum of %d "
                        "(B2RT_COUNT_HITS_STATIC_MAX); use b2rt_count_hits or split the batch", (long long)num_rays, B2RT_COUNT_HITS_STATIC_MAX);
    unsigned long long* scratch = nullptr;
    int rc = scratchForCurrentDevice(&scratch);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    cudaError_t e;
    if (num_rays == 0) {
        e = cudaMemsetAsync(hits_dev, 0, sizeof(int64_t), s);
        if (e != cudaSuccess) return setError(nullptr, B2RT_ERR_CUDA, "b2rt_count_hits_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
        return B2RT_OK;
    }
    e = cudaMemsetAsync(scratch, 0, sizeof(unsigned long long), s);          // the ticket starts at zero whatever the last call left
    if (e != cudaSuccess) return setError(nullptr, B2RT_ERR_CUDA, "b2rt_count_hits_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
    const unsigned grid = (unsigned)((num_rays + kStaticTile - 1) / kStaticTile);
    countHitsStaticKernel<<<grid, 256, 0, s>>>(static_cast<const int4*>(results_dev), (long long)num_rays, scratch,
                                               reinterpret_cast<long long*>(hits_dev));
    e = cudaGetLastError();
    if (e != cudaSuccess) return setError(nullptr, B2RT_ERR_CUDA, "launch of countHitsStaticKernel failed: %s", cudaGetErrorString(e));
    return B2RT_OK;
}
