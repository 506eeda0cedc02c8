// raysort.cu -- Morton-code ray sort, entirely on the device.
//
// Replaces RayBuffer::mortonSort (ray/RayBuffer.cc:256-324), findAABBKernel /
// genMortonKeysKernel / reorderRaysKernel (ray/RayBufferKernels.cu:74-200) and, above all, the
// D2H copy -> single-thread CPU quicksort -> H2D copy in the middle of it (RayBuffer.cc:291-297,
// framework/base/Sort.cc:131-165).
//
//  1. AABB of ray origins and end points: grid-stride loop, warp shuffles, one ordered-uint
//     atomicMin/Max per block (the reference reduces through shared memory without __syncwarp
//     and finishes with a float CAS loop, RayBufferKernels.cu:38-62,108-113).
//  2. 192-bit key: bit i of component c lands at key bit 6i + c (collectBits, :66-70); done with
//     compile-time shifts in registers, no shared-memory read-modify-writes.  Stored as three
//     64-bit words (lo/mid/hi) for the sort and, for b2rt_morton_keys, as the reference's 28-byte
//     MortonKey records.
//  3. Sort by a PREFIX of the key, then settle the ties.  Keys of in-range rays end at bit 146 (origin bit 24 at 6*24+2);
//     the prefix is key >> 91 (56 bits: the top 9-10 bits of every origin component, the top 6-7 of every direction
//     component), saturated -- a monotone function of the key, so sorting by (prefix, key) is sorting by the key.
//     Measured on a 2 Mi-ray AO batch: shift 83 (8 passes) 0.452 ms, 91 (7) 0.425, 99 (6) 0.452 -- fewer passes, more and
//     longer ties -- 107 falls back.  (a) stable LSD radix sort of (prefix, slot) pairs, 8 bits per pass, 7 passes: block
//     histograms, one scan kernel (one CTA per bin), a stable scatter that ranks with ballots -- no atomics on the data
//     path; (b) one thread per run of equal prefixes orders the run's slots by the full 192-bit key (stable insertion
//     sort; the words are gathered through the slots).  A run longer than kMaxRun (thousands of rays with one origin and
//     direction cell) raises a flag and the call falls back to (c): the full LSD sort over all three 64-bit words -- 8 + 8
//     passes and as many as the high word has significant bytes -- which is what every call did before (20 passes,
//     0.82 ms per 2 Mi rays).
//  4. Gather of rays by sorted slot + rebuild of idToSlot / slotToID (reorderRaysKernel).
// Compiled with -fmad=false so the quantised components are bit-identical to the CPU oracle.
#include "common.cuh"

namespace b2rt {

typedef unsigned long long u64;

__device__ __forceinline__ unsigned orderedBits(float f)
{
    const unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float orderedToFloat(unsigned u)
{
    const unsigned b = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    float f;
    memcpy(&f, &b, 4);
    return f;
}

__global__ void aabbInitKernel(unsigned* box)
{
    if (threadIdx.x < 3) box[threadIdx.x] = 0xffffffffu;        // lo: identity of min
    else if (threadIdx.x < 6) box[threadIdx.x] = 0u;            // hi: identity of max
}

__global__ void __launch_bounds__(256) findAABBKernel(const float4* __restrict__ rays, int n, unsigned* box)
{
    const float big = 3.402823466e+38f;
    float lo[3] = { big, big, big }, hi[3] = { -big, -big, -big };
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float4 o = __ldg(rays + 2 * i), d = __ldg(rays + 2 * i + 1);
        const float e[3] = { o.x + d.x * d.w, o.y + d.y * d.w, o.z + d.z * d.w };
        const float oo[3] = { o.x, o.y, o.z };
#pragma unroll
        for (int k = 0; k < 3; k++) {
            lo[k] = fminf(fminf(lo[k], oo[k]), e[k]);
            hi[k] = fmaxf(fmaxf(hi[k], oo[k]), e[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++)
        for (int s = 16; s > 0; s >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], s));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], s));
        }
    __shared__ float s_lo[8][3], s_hi[8][3];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0)
        for (int k = 0; k < 3; k++) { s_lo[warp][k] = lo[k]; s_hi[warp][k] = hi[k]; }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int k = threadIdx.x;
        float a = s_lo[0][k], b = s_hi[0][k];
        for (int w = 1; w < (int)(blockDim.x >> 5); w++) { a = fminf(a, s_lo[w][k]); b = fmaxf(b, s_hi[w][k]); }
        atomicMin(box + k, orderedBits(a));
        atomicMax(box + 3 + k, orderedBits(b));
    }
}

// hash[(6i+c) >> 5] |= bit i of v << ((6i+c) & 31), for i = 0..31 -- all shifts are constants
template <int C>
__device__ __forceinline__ void scatterBits(unsigned (&hash)[6], unsigned v)
{
#pragma unroll
    for (int i = 0; i < 32; i++) {
        const int pos = 6 * i + C, w = pos >> 5, sh = pos & 31;
        const unsigned moved = (sh >= i) ? (v << (sh - i)) : (v >> (i - sh));
        hash[w] |= moved & (1u << sh);
    }
}

// Prefix of the 192-bit key the first stage sorts by: bits shift .. 146, prefixPasses(shift) bytes wide.
constexpr int kKeyTopBit = 146;
constexpr int kPrefixShiftDefault = 91;                                     // 7 passes (64 <= shift <= 139: 8 .. 1 passes)
constexpr int kMaxRun = 48;                                                 // longest run of equal prefixes settled by one thread
static inline int prefixPasses(int shift) { return (kKeyTopBit + 1 - shift + 7) / 8; }

__device__ __forceinline__ u64 keyPrefix(u64 mid, u64 hi, int shift, int passes)
{
    const u64 all = (passes == 8) ? ~0ull : ((1ull << (8 * passes)) - 1ull);
    if (hi >> (kKeyTopBit + 1 - 128)) return all;          // out-of-range ray (bits above 146): after every in-range key
    return ((mid >> (shift - 64)) | (hi << (128 - shift))) & all;
}

// box: ordered-uint AABB (device) ; boxF: optional float AABB (for b2rt_morton_keys)
template <bool AOS>
__global__ void __launch_bounds__(256) mortonKeysKernel(const float4* __restrict__ rays, int n, const unsigned* box,
                                                        const float* boxF, u64* lo, u64* mid, u64* hi, u64* prefix,
                                                        unsigned* slots, MortonKey* keysAos, int prefixShift, int prefixBytes,
                                                        float4* rayCopy, const int32_t* slotToID, int32_t* slotToIDCopy)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float bl[3], bh[3];
    if (boxF) { for (int k = 0; k < 3; k++) { bl[k] = boxF[k]; bh[k] = boxF[3 + k]; } }
    else      { for (int k = 0; k < 3; k++) { bl[k] = orderedToFloat(box[k]); bh[k] = orderedToFloat(box[3 + k]); } }
    const float4 o = __ldg(rays + 2 * (size_t)i), d = __ldg(rays + 2 * (size_t)i + 1);
    const float ax = (o.x - bl[0]) / (bh[0] - bl[0]);
    const float ay = (o.y - bl[1]) / (bh[1] - bl[1]);
    const float az = (o.z - bl[2]) / (bh[2] - bl[2]);
    float l = 0.0f;
    l += d.x * d.x; l += d.y * d.y; l += d.z * d.z;
    const float len = sqrtf(l);
    const float s = 1.0f * ((len != 0.0f) ? 1.0f / len : 0.0f);
    const float bx = (d.x * s + 1.0f) * 0.5f, by = (d.y * s + 1.0f) * 0.5f, bz = (d.z * s + 1.0f) * 0.5f;
    unsigned hash[6] = { 0, 0, 0, 0, 0, 0 };
    // (U32)f on the device saturates: NaN and negatives -> 0, too large -> 0xffffffff
    scatterBits<0>(hash, __float2uint_rz(ax * 256.0f * 65536.0f));
    scatterBits<1>(hash, __float2uint_rz(ay * 256.0f * 65536.0f));
    scatterBits<2>(hash, __float2uint_rz(az * 256.0f * 65536.0f));
    scatterBits<3>(hash, __float2uint_rz(bx * 32.0f * 65536.0f));
    scatterBits<4>(hash, __float2uint_rz(by * 32.0f * 65536.0f));
    scatterBits<5>(hash, __float2uint_rz(bz * 32.0f * 65536.0f));
    if (AOS) {
        MortonKey k;
        k.oldSlot = i;
        for (int j = 0; j < 6; j++) k.hash[j] = hash[j];
        keysAos[i] = k;
    } else {
        lo[i] = (u64)hash[0] | ((u64)hash[1] << 32);
        mid[i] = (u64)hash[2] | ((u64)hash[3] << 32);
        hi[i] = (u64)hash[4] | ((u64)hash[5] << 32);
        // the sort is in place: the gather at its end reads these copies (made here, where the rays pass through registers
        // anyway, instead of by two device-to-device copies afterwards)
        rayCopy[2 * (size_t)i] = o; rayCopy[2 * (size_t)i + 1] = d;
        slotToIDCopy[i] = slotToID[i];
        prefix[i] = keyPrefix((u64)hash[2] | ((u64)hash[3] << 32), (u64)hash[4] | ((u64)hash[5] << 32), prefixShift, prefixBytes);
        slots[i] = (unsigned)i;
    }
}

// ---- radix sort ------------------------------------------------------------------------------
// 2 048-key tiles (256 threads x 8 keys in both kernels).  4 096-key tiles on 512 scatter threads were 4 % slower end to end: at 64
// registers an SM holds 1 024 scatter threads either way, but small blocks refill one at a time as they finish, so the partial
// last wave costs a short tile, not a long one (1 536-key tiles: slower again, the per-digit write runs get too short).
constexpr int kRsThreads = 256, kRsItems = 8, kRsTile = kRsThreads * kRsItems, kRsBins = 256;    // histogram kernel
// scatter kernel: the same tile, 8 keys per thread
constexpr int kScThreads = 256, kScItems = kRsTile / kScThreads, kScWarps = kScThreads / 32, kScWarpTile = kRsTile / kScWarps;

// blockHist layout: [bin][block]
__global__ void __launch_bounds__(kRsThreads) rsHistKernel(const u64* __restrict__ keys, int n, int shift,
                                                           unsigned* blockHist, int numBlocks, u64* orReduce)
{
    __shared__ unsigned hist[kRsBins];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const long long base = (long long)blockIdx.x * kRsTile;
    u64 ored = 0;
    // all 16 loads of the thread first (independent, in flight together), then the counting: with the loads issued four at a
    // time the kernel was latency-bound, 11 us for 16 MB
    u64 k[kRsItems];
#pragma unroll
    for (int j = 0; j < kRsItems; j++) {
        const long long i = base + j * kRsThreads + threadIdx.x;
        k[j] = (i < n) ? keys[i] : 0ull;
    }
#pragma unroll
    for (int j = 0; j < kRsItems; j++) {
        const long long i = base + j * kRsThreads + threadIdx.x;
        if (i < n) {
            ored |= k[j];
            atomicAdd(&hist[(unsigned)(k[j] >> shift) & 0xffu], 1u);
        }
    }
    __syncthreads();
    blockHist[(size_t)threadIdx.x * numBlocks + blockIdx.x] = hist[threadIdx.x];
    if (orReduce) {     // OR of all keys of this word: tells the host-free pass loop which digits are all zero
        for (int s = 16; s > 0; s >>= 1) ored |= __shfl_xor_sync(0xffffffffu, ored, s);
        if ((threadIdx.x & 31) == 0 && ored) atomicOr(orReduce, ored);
    }
}

// one CTA per bin: exclusive scan of that bin's row, total to binTotal[bin]
__global__ void __launch_bounds__(512) rsScanKernel(unsigned* blockHist, int numBlocks, unsigned* binTotal)
{
    __shared__ unsigned warpSum[16];
    __shared__ unsigned carry;
    unsigned* row = blockHist + (size_t)blockIdx.x * numBlocks;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < numBlocks; base += 512) {
        const int i = base + threadIdx.x;
        const unsigned v = (i < numBlocks) ? row[i] : 0u;
        unsigned incl = v;
        for (int s = 1; s < 32; s <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += t; }
        if (lane == 31) warpSum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            unsigned w = (lane < 16) ? warpSum[lane] : 0u;
            for (int s = 1; s < 16; s <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, w, s); if (lane >= s) w += t; }
            if (lane < 16) warpSum[lane] = w;       // inclusive over warps
        }
        __syncthreads();
        const unsigned warpBase = (warp == 0) ? 0u : warpSum[warp - 1];
        const unsigned c = carry;
        if (i < numBlocks) row[i] = c + warpBase + incl - v;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + warpSum[15];
        __syncthreads();
    }
    if (threadIdx.x == 0) binTotal[blockIdx.x] = carry;
}

// Stable scatter of one tile.  The tile is first put in digit order in shared memory (rank = start of the digit inside the
// tile + same-digit items of earlier warps + same-digit items earlier in this warp), then written out linearly: consecutive
// threads write consecutive elements of a digit's run, so the global stores are whole lines (2 048 items over 256 digits =
// runs of 8 keys = 64 B on average, two sectors) instead of one 32-byte sector per key.  (Scattering straight from registers ran at 0.96 TB/s
// of useful traffic, 50 us per pass on 2 Mi keys.)
// Lanes of the warp holding the same digit as this one, from nine ballots (one per digit bit + validity).  match.any does
// the same in one instruction but at a fraction of the throughput: with it the ranking loops were most of a pass.
__device__ __forceinline__ unsigned sameDigitLanes(unsigned digit, bool valid)
{
    unsigned peers = __ballot_sync(0xffffffffu, valid);
    if (!valid) peers = ~peers;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const bool bit = (digit >> b) & 1u;
        const unsigned bal = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
}

__global__ void __launch_bounds__(kScThreads) rsScatterKernel(const u64* __restrict__ keysIn, const unsigned* __restrict__ slotsIn,
                                                              u64* keysOut, unsigned* slotsOut, int n, int shift,
                                                              const unsigned* __restrict__ blockHist, int numBlocks,
                                                              const unsigned* __restrict__ binTotal)
{
    __shared__ unsigned warpHist[kScWarps][kRsBins];
    __shared__ unsigned globalBase[kRsBins];            // destination of the digit's first item of this tile, minus its place in the tile
    __shared__ unsigned scanTmp[8];
    extern __shared__ u64 s_tile[];                     // kRsTile keys, then kRsTile slots (24 KB)
    u64* sKey = s_tile;
    unsigned* sSlot = reinterpret_cast<unsigned*>(s_tile + kRsTile);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool binThread = threadIdx.x < kRsBins;       // threads 0..255 (warps 0..7) also act as "one thread per bin"
    for (int i = threadIdx.x; i < kScWarps * kRsBins; i += kScThreads) (&warpHist[0][0])[i] = 0;

    // all of this thread's keys and slots first: independent loads, in flight together
    const long long tileBase = (long long)blockIdx.x * kRsTile;
    const long long warpBase = tileBase + (long long)warp * kScWarpTile;
    u64 key[kScItems];
    unsigned slot[kScItems], digit[kScItems], peersOf[kScItems];
#pragma unroll
    for (int j = 0; j < kScItems; j++) {
        const long long i = warpBase + j * 32 + lane;
        const bool valid = i < n;
        key[j] = valid ? keysIn[i] : 0ull;
        slot[j] = valid ? slotsIn[i] : 0u;
        digit[j] = valid ? ((unsigned)(key[j] >> shift) & 0xffu) : 0xffffffffu;
    }
    // exclusive scan of the 256 bin totals
    unsigned binBase = 0;
    {
        unsigned v = 0, incl = 0;
        if (binThread) {
            v = binTotal[threadIdx.x];
            incl = v;
            for (int s = 1; s < 32; s <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += t; }
            if (lane == 31) scanTmp[warp] = incl;
        }
        __syncthreads();                                // also: warpHist is zero
        if (binThread) {
            unsigned wb = 0;
            for (int w = 0; w < warp; w++) wb += scanTmp[w];
            binBase = wb + incl - v;
        }
    }
    // phase A: per-warp digit counts, in element order
#pragma unroll
    for (int j = 0; j < kScItems; j++) {
        const bool valid = digit[j] != 0xffffffffu;
        const unsigned peers = sameDigitLanes(digit[j], valid);
        peersOf[j] = peers;
        if (valid && lane == __ffs(peers) - 1) warpHist[warp][digit[j]] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // phase B: counts -> places inside the tile.  Thread = bin: tile total of the bin, exclusive scan over the bins, then
    // the start of every warp's share of the bin.
    {
        const int bin = threadIdx.x;
        unsigned total = 0, incl = 0;
        if (binThread) {
            for (int w = 0; w < kScWarps; w++) total += warpHist[w][bin];
            incl = total;
            for (int s = 1; s < 32; s <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += t; }
            if (lane == 31) scanTmp[warp] = incl;       // (the first scan's readers are past the barrier above)
        }
        __syncthreads();
        if (binThread) {
            unsigned wb = 0;
            for (int w = 0; w < warp; w++) wb += scanTmp[w];
            const unsigned tileStart = wb + incl - total;   // first place of this bin inside the tile
            globalBase[bin] = binBase + blockHist[(size_t)bin * numBlocks + blockIdx.x] - tileStart;
            unsigned running = tileStart;
            for (int w = 0; w < kScWarps; w++) { const unsigned c = warpHist[w][bin]; warpHist[w][bin] = running; running += c; }
        }
    }
    __syncthreads();
    // phase C: stable placement in shared memory
#pragma unroll
    for (int j = 0; j < kScItems; j++) {
        const bool valid = digit[j] != 0xffffffffu;
        const unsigned peers = peersOf[j];
        if (valid) {
            const unsigned place = warpHist[warp][digit[j]] + __popc(peers & ((1u << lane) - 1u));
            sKey[place] = key[j];
            sSlot[place] = slot[j];
        }
        __syncwarp();
        if (valid && lane == __ffs(peers) - 1) warpHist[warp][digit[j]] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // phase D: linear write-out
    const int count = (int)min((long long)kRsTile, (long long)n - tileBase);
#pragma unroll
    for (int j = 0; j < kScItems; j++) {
        const int place = j * kScThreads + threadIdx.x;
        if (place < count) {
            const u64 k = sKey[place];
            const unsigned dst = globalBase[(unsigned)(k >> shift) & 0xffu] + place;
            keysOut[dst] = k;
            slotsOut[dst] = sSlot[place];
        }
    }
}

// Stage (b): runs of equal prefixes.  The thread at the head of a run of 2..kMaxRun elements orders the run's slots by the
// full key (hi, mid, lo words gathered through the slots); the prefix sort was stable, so the run arrives in slot order and
// a stable insertion sort leaves equal keys in slot order.  Longer runs set *tooLong (the caller then sorts the long way).
__global__ void __launch_bounds__(256) settleRunsKernel(const u64* __restrict__ prefix, unsigned* slots, int n,
                                                        const u64* __restrict__ lo, const u64* __restrict__ mid,
                                                        const u64* __restrict__ hi, int* tooLong)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 p = prefix[i];
    if (i > 0 && prefix[i - 1] == p) return;            // not the head of its run
    if (i + 1 >= n || prefix[i + 1] != p) return;       // a run of one
    int len = 2;
    while (len <= kMaxRun && i + len < n && prefix[i + len] == p) len++;
    if (len > kMaxRun) { *tooLong = 1; return; }
    unsigned sl[kMaxRun];
    u64 kh[kMaxRun], km[kMaxRun], kl[kMaxRun];
    for (int j = 0; j < len; j++) {
        const unsigned s = slots[i + j];
        const u64 h = hi[s], m = mid[s], l = lo[s];
        int k = j;
        while (k > 0 && (kh[k - 1] > h || (kh[k - 1] == h && (km[k - 1] > m || (km[k - 1] == m && kl[k - 1] > l))))) {
            kh[k] = kh[k - 1]; km[k] = km[k - 1]; kl[k] = kl[k - 1]; sl[k] = sl[k - 1];
            k--;
        }
        kh[k] = h; km[k] = m; kl[k] = l; sl[k] = s;
    }
    for (int j = 0; j < len; j++) slots[i + j] = sl[j];
}

__global__ void __launch_bounds__(256) iotaKernel(unsigned* slots, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) slots[i] = (unsigned)i;
}

__global__ void __launch_bounds__(256) gatherWordKernel(const u64* __restrict__ word, const unsigned* __restrict__ slots, u64* out, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = word[slots[i]];
}

// reorderRaysKernel -- RayBufferKernels.cu:183-200
__global__ void __launch_bounds__(256) reorderRaysKernel(const unsigned* __restrict__ sortedSlots, const float4* __restrict__ inRays,
                                                         const int32_t* __restrict__ inSlotToID, float4* outRays,
                                                         int32_t* outIDToSlot, int32_t* outSlotToID, int n, int slotStride)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned oldSlot = sortedSlots[(size_t)i * slotStride];   // stride 7: oldSlot of 28-byte MortonKey records
    const int id = inSlotToID[oldSlot];
    outRays[2 * (size_t)i] = __ldg(inRays + 2 * (size_t)oldSlot);
    outRays[2 * (size_t)i + 1] = __ldg(inRays + 2 * (size_t)oldSlot + 1);
    outIDToSlot[id] = i;
    outSlotToID[i] = id;
}

static inline size_t alignUp(size_t v) { return (v + 255) & ~(size_t)255; }

static int findAABB(b2rt_ctx* ctx, const void* rays, int n, cudaStream_t s)
{
    unsigned* box = reinterpret_cast<unsigned*>(ctx->aabbDev);
    aabbInitKernel<<<1, 32, 0, s>>>(box);
    B2RT_CHECK_LAUNCH(ctx, "aabbInitKernel");
    int grid = (n + 255) / 256;
    const int cap = ctx->smCount * 8;
    if (grid > cap) grid = cap;
    findAABBKernel<<<grid, 256, 0, s>>>(static_cast<const float4*>(rays), n, box);
    B2RT_CHECK_LAUNCH(ctx, "findAABBKernel");
    return B2RT_OK;
}

static int readAABB(b2rt_ctx* ctx, float out[6], cudaStream_t s)
{
    unsigned h[6];
    B2RT_CUDA(ctx, cudaMemcpyAsync(h, ctx->aabbDev, sizeof(h), cudaMemcpyDeviceToHost, s));
    B2RT_CUDA(ctx, cudaStreamSynchronize(s));
    for (int k = 0; k < 6; k++) out[k] = orderedToFloat(h[k]);
    return B2RT_OK;
}

}  // namespace b2rt

using namespace b2rt;

extern "C" {

int b2rt_find_aabb(b2rt_ctx* ctx, const void* rays, int64_t n, float aabbOut[6], void* stream)
{
    if (!ctx || !aabbOut) return B2RT_ERR_ARG;
    if (n < 0 || n > 0x7fffffff || (n && !rays)) return setError(ctx, B2RT_ERR_ARG, "b2rt_find_aabb: bad arguments");
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    int rc = findAABB(ctx, rays, (int)n, s);
    if (rc) return rc;
    return readAABB(ctx, aabbOut, s);
}

int b2rt_morton_keys(b2rt_ctx* ctx, const void* rays, int64_t n, const float aabb[6], void* keys, void* stream)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (n < 0 || n > 0x7fffffff || !aabb || (n && (!rays || !keys))) return setError(ctx, B2RT_ERR_ARG, "b2rt_morton_keys: bad arguments");
    if (n == 0) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    float* boxF = ctx->aabbDev;     // 8 floats allocated; reuse as float box
    B2RT_CUDA(ctx, cudaMemcpyAsync(boxF, aabb, 6 * sizeof(float), cudaMemcpyHostToDevice, s));
    mortonKeysKernel<true><<<(unsigned)((n + 255) / 256), 256, 0, s>>>(static_cast<const float4*>(rays), (int)n, nullptr, boxF,
                                                                      nullptr, nullptr, nullptr, nullptr, nullptr, static_cast<MortonKey*>(keys), 0, 0,
                                                                      nullptr, nullptr, nullptr);
    B2RT_CHECK_LAUNCH(ctx, "mortonKeysKernel");
    B2RT_CUDA(ctx, cudaStreamSynchronize(s));   // aabb is a host pointer the caller may reuse
    return B2RT_OK;
}

int b2rt_morton_sort(b2rt_ctx* ctx, void* rays, int32_t* idToSlot, int32_t* slotToID, int64_t n64, float aabbOut[6], void* stream)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (n64 < 0 || n64 > 0x7fffffff || (n64 && (!rays || !idToSlot || !slotToID)))
        return setError(ctx, B2RT_ERR_ARG, "b2rt_morton_sort: bad arguments");
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int n = (int)n64;
    if (n == 0) {
        if (aabbOut) { for (int k = 0; k < 3; k++) { aabbOut[k] = 3.402823466e+38f; aabbOut[3 + k] = -3.402823466e+38f; } }
        return B2RT_OK;
    }
    const int numBlocks = (n + kRsTile - 1) / kRsTile;
    // scratch layout
    const size_t wordB = alignUp((size_t)n * 8), slotB = alignUp((size_t)n * 4);
    const size_t histB = alignUp((size_t)kRsBins * numBlocks * 4);
    const size_t rayB = alignUp((size_t)n * 32);
    size_t total = 3 * wordB + 2 * wordB + 2 * slotB + histB + 4096 + rayB + slotB;
    int rc = ensureScratch(ctx, total);
    if (rc) return rc;
    char* base = static_cast<char*>(ctx->scratch);
    u64* wLo = reinterpret_cast<u64*>(base); base += wordB;
    u64* wMid = reinterpret_cast<u64*>(base); base += wordB;
    u64* wHi = reinterpret_cast<u64*>(base); base += wordB;
    u64* kA = reinterpret_cast<u64*>(base); base += wordB;
    u64* kB = reinterpret_cast<u64*>(base); base += wordB;
    unsigned* sA = reinterpret_cast<unsigned*>(base); base += slotB;
    unsigned* sB = reinterpret_cast<unsigned*>(base); base += slotB;
    unsigned* blockHist = reinterpret_cast<unsigned*>(base); base += histB;
    unsigned* binTotal = reinterpret_cast<unsigned*>(base);
    u64* orWord = reinterpret_cast<u64*>(base + 2048);
    int* flag = reinterpret_cast<int*>(base + 3072); base += 4096;
    float4* oldRays = reinterpret_cast<float4*>(base); base += rayB;
    int32_t* oldSlotToID = reinterpret_cast<int32_t*>(base);

    B2RT_CUDA(ctx, cudaFuncSetAttribute(rsScatterKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRsTile * 12));
    rc = findAABB(ctx, rays, n, s);
    if (rc) return rc;
    const unsigned g = (unsigned)((n + 255) / 256);
    const int prefixShift = ctx->sortMode >= 64 ? ctx->sortMode : kPrefixShiftDefault;
    const int passesA = prefixPasses(prefixShift);
    mortonKeysKernel<false><<<g, 256, 0, s>>>(static_cast<const float4*>(rays), n, reinterpret_cast<unsigned*>(ctx->aabbDev), nullptr,
                                             wLo, wMid, wHi, kA, sA, nullptr, prefixShift, passesA, oldRays, slotToID, oldSlotToID);
    B2RT_CHECK_LAUNCH(ctx, "mortonKeysKernel");

    unsigned* slotsCur = sA; unsigned* slotsAlt = sB;
    // one LSD pass over byte `pass` of the words in `cur`; returns the buffer the pass wrote
    auto radixPass = [&](const u64* cur, int pass, bool histTaken) -> const u64* {
        u64* outBuf = (cur == kA) ? kB : kA;
        if (!histTaken) rsHistKernel<<<numBlocks, kRsThreads, 0, s>>>(cur, n, pass * 8, blockHist, numBlocks, nullptr);
        rsScanKernel<<<kRsBins, 512, 0, s>>>(blockHist, numBlocks, binTotal);
        rsScatterKernel<<<numBlocks, kScThreads, kRsTile * 12, s>>>(cur, slotsCur, outBuf, slotsAlt, n, pass * 8, blockHist, numBlocks, binTotal);
        ctx->launches += histTaken ? 2 : 3;
        unsigned* t = slotsCur; slotsCur = slotsAlt; slotsAlt = t;
        return outBuf;
    };

    // (a) by prefix, (b) ties by the full key
    int tooLong = 0;
    if (ctx->sortMode != 1) {
        const u64* cur = kA;
        for (int pass = 0; pass < passesA; pass++) cur = radixPass(cur, pass, false);
        B2RT_CUDA(ctx, cudaGetLastError());
        B2RT_CUDA(ctx, cudaMemsetAsync(flag, 0, sizeof(int), s));
        settleRunsKernel<<<g, 256, 0, s>>>(cur, slotsCur, n, wLo, wMid, wHi, flag);
        B2RT_CHECK_LAUNCH(ctx, "settleRunsKernel");
        B2RT_CUDA(ctx, cudaMemcpyAsync(&tooLong, flag, sizeof(int), cudaMemcpyDeviceToHost, s));
        B2RT_CUDA(ctx, cudaStreamSynchronize(s));
    }
    ctx->sortFellBack = (ctx->sortMode == 1 || tooLong) ? 1 : 0;
    if (ctx->sortFellBack) {
        // (c) the long way: LSD over the whole key, word by word.  Keys of rays inside the AABB end at bit 146, so the high
        // word normally needs 3 passes; out-of-range input (NaN boxes etc.) may set more bits, so its pass count comes from
        // an OR-reduction of the word.
        if (slotsCur != sA) { unsigned* t = slotsCur; slotsCur = slotsAlt; slotsAlt = t; }
        iotaKernel<<<g, 256, 0, s>>>(slotsCur, n);
        B2RT_CHECK_LAUNCH(ctx, "iotaKernel");
        const u64* words[3] = { wLo, wMid, wHi };
        for (int w = 0; w < 3; w++) {
            const u64* cur = wLo;
            if (w > 0) {
                gatherWordKernel<<<g, 256, 0, s>>>(words[w], slotsCur, kA, n);
                B2RT_CHECK_LAUNCH(ctx, "gatherWordKernel");
                cur = kA;
            }
            int passes = 8;
            if (w == 2) {
                B2RT_CUDA(ctx, cudaMemsetAsync(orWord, 0, 8, s));
                rsHistKernel<<<numBlocks, kRsThreads, 0, s>>>(cur, n, 0, blockHist, numBlocks, orWord);
                B2RT_CHECK_LAUNCH(ctx, "rsHistKernel");
                u64 h = 0;
                B2RT_CUDA(ctx, cudaMemcpyAsync(&h, orWord, 8, cudaMemcpyDeviceToHost, s));
                B2RT_CUDA(ctx, cudaStreamSynchronize(s));
                passes = 0;
                while (h) { passes++; h >>= 8; }
            }
            for (int pass = 0; pass < passes; pass++) cur = radixPass(cur, pass, w == 2 && pass == 0);
            B2RT_CUDA(ctx, cudaGetLastError());
        }
    }

    reorderRaysKernel<<<g, 256, 0, s>>>(slotsCur, oldRays, oldSlotToID, static_cast<float4*>(rays), idToSlot, slotToID, n, 1);
    B2RT_CHECK_LAUNCH(ctx, "reorderRaysKernel");
    if (aabbOut) return readAABB(ctx, aabbOut, s);
    return B2RT_OK;
}

int b2rt_set_sort_mode(b2rt_ctx* ctx, int mode)
{
    if (!ctx || mode < 0 || (mode > 1 && (mode < 64 || mode > 139)))
        return setError(ctx, B2RT_ERR_ARG, "b2rt_set_sort_mode: mode must be 0, 1 or a prefix shift in 64..139");
    ctx->sortMode = mode;
    return B2RT_OK;
}

int b2rt_sort_fell_back(const b2rt_ctx* ctx) { return ctx ? ctx->sortFellBack : 0; }

int b2rt_reorder_rays(b2rt_ctx* ctx, const void* keys, const void* inRays, const int32_t* inSlotToID, void* outRays,
                      int32_t* outIDToSlot, int32_t* outSlotToID, int64_t n, void* stream)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (n < 0 || n > 0x7fffffff || (n && (!keys || !inRays || !inSlotToID || !outRays || !outIDToSlot || !outSlotToID)) ||
        inRays == outRays)
        return setError(ctx, B2RT_ERR_ARG, "b2rt_reorder_rays: bad arguments");
    if (n == 0) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    reorderRaysKernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(static_cast<const unsigned*>(keys), static_cast<const float4*>(inRays),
                                                                 inSlotToID, static_cast<float4*>(outRays), outIDToSlot, outSlotToID,
                                                                 (int)n, 7);
    B2RT_CHECK_LAUNCH(ctx, "reorderRaysKernel");
    return B2RT_OK;
}

}  // extern "C"
