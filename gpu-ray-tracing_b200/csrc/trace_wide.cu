// trace_wide.cu -- closest-hit / any-hit traversal of the PRIVATE 8-wide compressed form of the
// caller's Compact2 BVH (wide_bvh.h; built by host/wide_bvh.cpp inside the context).
//
// Why it exists: ncu on the binary kernel (profiles/r01_ncu_trace_c2.md) shows the LSU data pipe at
// 87 % of peak while active and `long_scoreboard` as the top stall -- every node visit costs each lane
// four 16-byte loads from a different 64-byte record.  An 8-wide node with 8-bit child boxes is 80
// bytes for EIGHT boxes, and a ray visits ~0.4x as many of them, so both the bytes through L1 and the
// number of dependent fetches per ray drop by more than half.  Same contract as trace.cu: persistent
// warps with a ballot-aggregated atomic ray fetch and dynamic refill, Woop test on the caller's own
// rows (so t, u, v are bit-identical to the binary path), one result store per ray.
//
// Traversal state per lane (after Ylitie et al., HPG 2017):
//   G = (base index of a group of sibling nodes, hits in bits 31..24 | imask in bits 7..0)
//   T = (base index of a group of triangles,     hits in bits 23..0)
// A node step pops the highest hit bit of G -- bit position = 24 + (slot XOR ray octant), so siblings
// come out front to back -- pushes the rest of G, tests the eight child boxes of that node and forms
// the next G and T from the resulting 32-bit hit mask.
//
// Shape of the shipped kernels (variants 108 / 114, launchTraceWide below; every choice with its measurement next to the code and in
// DESIGN.md section 4): 256-thread CTAs, four per SM at 64 registers; 8 stack entries per lane in shared memory, the rest in local
// memory; the triangle loop requests row 0 of a group's next triangle before it tests the current one and is unrolled twice; a
// triangle group that fewer than a quarter of the warp's lanes could be tested with waits in a one-entry room per lane (PARK) while
// the lane traverses on; rays and records move with the streaming cache policy; for batches above 12 Mi rays a warp claims 128
// rays per atomic and classifies the block at once (SCAN), smaller batches take exactly as many rays as there are idle lanes.
#include "common.cuh"

namespace b2rt {

struct WideParams {
    const float4* __restrict__ rays;
    int4* __restrict__ results;
    const float4* __restrict__ nodes;       // 5 float4 per node
    const float4* __restrict__ woop;        // 3 float4 per triangle
    const int* __restrict__ triIndex;       // 1 per triangle
    int numRays;
    int fetchThreshold;
    int pushLimit;                          // a triangle group may be postponed only while fewer than this many entries are stacked
    int streamRays;                         // 1: rays are read and records written with the streaming (evict-first) policy
    int* warpCounter;
    unsigned long long* stats;
    unsigned one;                           // 0x3F800000, passed as data so PRMT keeps its selector as the immediate
};

// Stack capacity is kWideStack + SSLOTS entries.  Every level of the wide tree can leave one sibling group on the stack and
// those pushes cannot be refused, so triangle groups are postponed only below WideParams::pushLimit = capacity - depth - 1 and
// the launcher sends trees deeper than capacity - 2 to the binary kernel (trace.cu:launchTrace).
constexpr int kWideStack = 48;
// The streaming policy for rays and records is an A/B switch (B2RT_TUNE_NO_STREAM) only in -DB2RT_EXPLORE builds: as a run-time
// flag it doubles every ray load and record store of the refill path into two predicated forms.
#ifdef B2RT_EXPLORE
constexpr bool kAlwaysStream = false;
#else
constexpr bool kAlwaysStream = true;
#endif
constexpr unsigned kFullMask = 0xffffffffu;


// byte j of `word` as the float 1 + byte * 2^-15 (exact): one PRMT, no integer-to-float conversion
// `one` must hold 0x3F800000 in a register the compiler cannot see through: PRMT takes one immediate, and it
// should be the SELECTOR (otherwise four selector constants get re-materialised before every PRMT).
template <int J>
__device__ __forceinline__ float byteToUnitFloat(unsigned word, unsigned one)
{
    unsigned r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(word), "r"(one), "n"(0x7604 | (J << 4)));
    return __uint_as_float(r);
}

template <bool BARY>
__device__ __forceinline__ void storeResultT(const WideParams& p, unsigned rayidx, int id, float t, float u, float v)
{
    if (BARY) {
        const int4 r = make_int4(id, __float_as_int(t), __float_as_int(u), __float_as_int(v));
        if (kAlwaysStream || p.streamRays) __stcs(p.results + rayidx, r); else p.results[rayidx] = r;
    } else {
        int2* dst = reinterpret_cast<int2*>(p.results + rayidx);           // 8 of the record's 16 bytes (CudaTracerKernels.hh:197)
        const int2 r = make_int2(id, __float_as_int(t));
        if (kAlwaysStream || p.streamRays) __stcs(dst, r); else *dst = r;
    }
}

// Measured and removed (round 2, profiles/r02_explore_c5_ring.log, r02_ncu_trace_wide_c5_ring.md): rays dealt out statically in
// 16-ray chunks through a per-warp shared-memory ring filled by cp.async one refill ahead, the work counter only for the last
// eighth of the batch.  It does what it was built for -- long_scoreboard stalls per issue fall from 2.78 to 1.39: half of them
// are the refill's atomic -> ray-load chain -- but a chunk is 16 samples of ONE primary ray, so 52 % of the C5 chunks are
// entirely degenerate and a refill often fills no lane: 20 % more instructions, 40 % instead of 50 % of the warp slots
// occupied, 3 265 -> 2 380 Mrays/s.
template <bool ANY_HIT, bool BARY, bool STATS, int BLOCK, int MIN_BLOCKS, int POSTPONE, int LEAF, int SSLOTS, bool WT, int CLAIM, int OPT>
__global__ void __launch_bounds__(BLOCK, MIN_BLOCKS) traceWideKernel(const WideParams p)
{
    extern __shared__ uint2 s_wideStack[];
    // OPT bit 1 (MASK): a child's (bits, bit index) byte is pulled out with one PRMT and shifted with a wrapping shift: 4 instead
    //           of 5 integer operations per child.  C5 3 495 -> 3 524..3 530 Mrays/s (+0.9 %).
    // (OPT bit 0 was HOLD: a lane whose turned-away group is all it has kept it in registers instead of pushing and popping it
    //  every iteration, and requested a group's first row only after the vote.  Same traversal (TRV / IST identical), fewer stack
    //  operations, 48 more static instructions -- and C5 3 495 -> 3 366: removed.  profiles/r02_explore_c5_last_session.log)
    constexpr bool MASK = (OPT & 2) != 0;
    // OPT bit 2 (PARK, LEAF == 1 only): a triangle group the vote turns away goes to a one-entry waiting room of its own (shared
    //           memory) instead of the traversal stack, and is offered again at EVERY triangle phase in which the lane has no
    //           fresh group; the lane keeps traversing meanwhile.  (A group pushed on the stack is not seen again before everything
    //           above it is done; a group popped from the stack and turned away again makes the lane spin on it.)  Node-loop SIMD
    //           efficiency on C5 77 -> 84 % at the same threshold; with the threshold at 1/4: 83 % / 32 %, C5 3 530 -> 3 556..3 585, C4
    //           612 -> 624..628, C3 3 330 -> 3 379..3 438, an 8 Mi-ray shard of C5 3 213 -> 3 251 (same log).
    constexpr bool PARK = (OPT & 4) != 0 && LEAF == 1 && POSTPONE > 0;
    constexpr int kParkDepth = (OPT & 8) ? 2 : 1;       // OPT bit 3: two waiting-room entries per lane (10 M of the 15 M warp iterations of a C5
                                                        // launch still push one lane's group because its room is taken -- but the indexed room
                                                        // costs more than those pushes: C5 3 585 -> 3 480)
    uint2* s_park = s_wideStack + SSLOTS * BLOCK + threadIdx.x;
    int parked = 0;                                     // groups of this lane in the waiting room
    // OPT bit 5 (SCAN, CLAIM > 0 only): a claimed block of rays is classified at once; see the fetch
    constexpr bool SCAN = (OPT & 32) != 0 && CLAIM > 0;
    unsigned* s_live = reinterpret_cast<unsigned*>(s_wideStack + (SSLOTS + (PARK ? kParkDepth : 0)) * BLOCK) + (threadIdx.x >> 5) * (CLAIM > 0 ? CLAIM : 1);
    auto storeResult = [](const WideParams& q, unsigned rayidx, int id, float t, float u, float v) { storeResultT<BARY>(q, rayidx, id, t, u, v); };
    const int lane = threadIdx.x & 31;
    // Traversal stack: the first SSLOTS entries in shared memory (slot k of this thread at s_stack[k * BLOCK]: consecutive
    // threads, consecutive 8-byte words), the rest in local memory.  Entirely in local memory (SSLOTS == 0) the stack moved
    // 23 GB through L1 per 64 Mi-ray launch and its pops held 6 % of the stall samples (ncu source page, C5).
    uint2* s_stack = s_wideStack + threadIdx.x;
    uint2 lstack[kWideStack];
    auto push = [&](int& sp_, const uint2 v) {
        if (sp_ < SSLOTS) s_stack[sp_ * BLOCK] = v; else lstack[sp_ - SSLOTS] = v;
        sp_++;
    };
    auto pop = [&](int& sp_) -> uint2 {
        sp_--;
        return (sp_ < SSLOTS) ? s_stack[sp_ * BLOCK] : lstack[sp_ - SSLOTS];
    };
    const unsigned one = p.one;                         // 0x3F800000 from the parameter block: opaque to ptxas (see byteToUnitFloat)

    float origx = 0, origy = 0, origz = 0, dirx = 0, diry = 0, dirz = 0;
    float idirx = 0, idiry = 0, idirz = 0, tmin = 0, hitT = 0, hitU = 0, hitV = 0;
    float cullT = 0;                    // WT: hitT widened for the box tests (common.cuh:woopAccept)
    int hitIndex = -1, rayidx = -1, sp = 0;
    unsigned octinv4 = 0;
    uint2 G = make_uint2(0u, 0u), T = make_uint2(0u, 0u);
    bool active = false;
    constexpr int kClaim = CLAIM;           // rays a warp claims per atomic on the work counter (0: as many as it has idle lanes)
    bool poolEmpty = false;
    unsigned blkBase = 0, blkLeft = 0;          // CLAIM: the warp's claimed block of rays (warp-uniform)
    unsigned long long stNodeIters = 0, stNodeLanes = 0, stTriIters = 0, stTriLanes = 0, stFetches = 0;

    // a fetched ray becomes this lane's ray -- or, degenerate (tmax < tmin, RayGenKernels.cu:223: it can hit nothing), is answered on the spot
    auto takeRay = [&](unsigned idx, const float4 o, const float4 d) {
        if (d.w < o.w) { storeResult(p, idx, -1, d.w, 0.0f, 0.0f); return; }
        rayidx = (int)idx;
        origx = o.x; origy = o.y; origz = o.z; tmin = o.w;
        dirx = d.x; diry = d.y; dirz = d.z; hitT = d.w;
        if (WT) cullT = hitT;
        const float ooeps = 8.271806125530277e-25f;    // 2^-80 (kepler_dynamic_fetch.cu:134-137)
        idirx = rcp_approx(fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x));
        idiry = rcp_approx(fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y));
        idirz = rcp_approx(fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z));
        // bit set <=> the ray travels towards + on that axis; replicated into four bytes
        octinv4 = ((idirx < 0.0f ? 0u : 4u) | (idiry < 0.0f ? 0u : 2u) | (idirz < 0.0f ? 0u : 1u)) * 0x01010101u;
        G = make_uint2(0u, 0x80000000u);               // the root, as a group of one
        T = make_uint2(0u, 0u);
        sp = 0; hitIndex = -1; parked = 0;
        if (BARY) { hitU = 0.0f; hitV = 0.0f; }
        active = true;
    };

    while (true) {
        __syncwarp();
        // ---- store finished rays, refill (same scheme as trace.cu) ----------------------------
        if (!active && rayidx >= 0) {
            const int id = (hitIndex < 0) ? -1 : __ldg(p.triIndex + hitIndex);
            storeResult(p, (unsigned)rayidx, id, hitT, hitU, hitV);
            rayidx = -1;
        }
        for (;;) {
            const unsigned maskT = __ballot_sync(kFullMask, !active);
            if (maskT == 0 || poolEmpty) break;
            unsigned numT = __popc(maskT);
            unsigned base;
            if (CLAIM > 0) {
                // Block claims: one atomic on the work counter buys the warp `claim` consecutive rays; refills are served from the
                // warp's own block, and the block's second half is pulled into L1 while the first is traced.  The refill chain
                // "atomic (L2 round trip under contention) -> dependent ray load (DRAM: the batch streams through once)" held 13 % of
                // the stall samples at 3 % of the instructions (profiles/r02_ncu_trace_wide_c5.md); 4 of 5 refills now find both
                // the index and the ray at hand.  Measured on the C5 batch (profiles/r02_explore_c5_claims.log): 64 rays per claim +0.7 %,
                // 128 +2.4 %, 256 +2.0 %; with the refill threshold at 26 instead of 24 another +0.6 %.  A warp holds up to claim - 1
                // unstarted rays when the pool runs dry, which lengthens the tail of a frame-sized launch (C4 -14 %, C3 -10 %): used for
                // batches above 12 Mi rays only (launchTrace; 24 Mi before block classification).
                if (blkLeft == 0u) {
                    unsigned b = 0;
                    if (lane == 0) b = (unsigned)atomicAdd(p.warpCounter, kClaim);
                    b = __shfl_sync(kFullMask, b, 0);
                    if (STATS && lane == 0) stFetches++;
                    if (b >= (unsigned)p.numRays) { poolEmpty = true; break; }
                    if (SCAN) {
                        // The whole block is looked at once, all lanes wide: tmin and tmax of every ray (which also brings the rays'
                        // sectors into L1), degenerate rays (tmax < tmin: the samples of a primary ray that hit nothing, 52 % of the C5
                        // batch, in runs of 86) answered here, the indices of the others written to the warp's list.  Refills then
                        // hand out live rays only: no refill round that fills no lane.  C5 3 719 -> 3 785 Mrays/s, node-loop SIMD efficiency
                        // 82.6 -> 85.5 %; the refill threshold stays at 26 (22: 3 699, 24: 3 767, 28: 3 759, 30: 3 657).
                        const unsigned n = min((unsigned)kClaim, (unsigned)p.numRays - b);
                        unsigned live = 0;
                        float tmn[kClaim / 32 > 0 ? kClaim / 32 : 1], tmx[kClaim / 32 > 0 ? kClaim / 32 : 1];
#pragma unroll
                        for (int k = 0; k < kClaim / 32; k++) {             // all loads first: one round trip for the block, not four
                            const unsigned off = 32u * k + lane;
                            tmn[k] = 0.0f; tmx[k] = -1.0f;
                            if (off < n) {
                                const float* r = reinterpret_cast<const float*>(p.rays + 2 * (size_t)(b + off));
                                tmn[k] = __ldg(r + 3); tmx[k] = __ldg(r + 7);
                            }
                        }
#pragma unroll
                        for (int k = 0; k < kClaim / 32; k++) {
                            const unsigned off = 32u * k + lane;
                            const bool ok = !(tmx[k] < tmn[k]);
                            if (!ok && off < n) storeResult(p, b + off, -1, tmx[k], 0.0f, 0.0f);
                            const unsigned m = __ballot_sync(kFullMask, ok);
                            if (ok) s_live[live + __popc(m & ((1u << lane) - 1u))] = b + off;
                            live += __popc(m);
                        }
                        __syncwarp();
                        blkBase = 0u; blkLeft = live;
                        if (live == 0u) continue;                   // a block of misses only: claim the next one
                    } else {
                    blkBase = b;
                    blkLeft = min((unsigned)kClaim, (unsigned)p.numRays - b);
                    // rays [b + 32, b + claim): 32 bytes each, one 128-byte line per four rays
#pragma unroll
                    for (int k = 0; k < (kClaim - 32 + 127) / 128; k++) {
                        const unsigned r = b + 32u + 4u * (lane + 32u * k);
                        if (4 * (lane + 32 * k) < kClaim - 32 && r < (unsigned)p.numRays)
                            asm volatile("prefetch.global.L1 [%0];" :: "l"(p.rays + 2 * (size_t)r));
                    }
                    }
                }
                numT = min(numT, blkLeft);
                base = blkBase;
                blkBase += numT; blkLeft -= numT;
            } else {
            const int leader = __ffs(maskT) - 1;
            base = 0;
            if (lane == leader) base = (unsigned)atomicAdd(p.warpCounter, (int)numT);
            base = __shfl_sync(kFullMask, base, leader);
            if (STATS && lane == leader) stFetches++;
            poolEmpty = (base + numT >= (unsigned)p.numRays);
            }
            if (!active) {
                const unsigned rank = __popc(maskT & ((1u << lane) - 1u));
                const unsigned idx = (SCAN && rank < numT) ? s_live[base + rank] : base + rank;
                if (rank < numT && idx < (unsigned)p.numRays) {
                    float4 o, d;
                    if (kAlwaysStream || p.streamRays) { o = __ldcs(p.rays + 2 * (size_t)idx); d = __ldcs(p.rays + 2 * (size_t)idx + 1); }
                    else { o = ldg4(p.rays + 2 * (size_t)idx); d = ldg4(p.rays + 2 * (size_t)idx + 1); }
                    takeRay(idx, o, d);
                }
            }
            // A fetch that brought only degenerate rays is repeated at once (nobody is waiting).  Repeating it until the refill
            // threshold is met whenever SOME of the fetched rays were degenerate was measured and lost 11 % on the C5 batch
            // (6.27 M instead of 4.64 M fetches, each a round trip the live lanes sit through): profiles/r02_explore_c5_switches.log.
            if (__ballot_sync(kFullMask, active) != 0) break;
        }
        if (__ballot_sync(kFullMask, active) == 0) {
            if (poolEmpty) break;                                       // nothing in flight and the pool is dry
            continue;                                                   // only degenerate rays so far: draw again
        }
        const int threshold = poolEmpty ? 0 : p.fetchThreshold;

        // `if (active) for (;;)` with explicit breaks, not `while (active)`: the loop leaves only through the breaks below (each sets what it
        // has to), and with the flag out of the loop condition ptxas drops 39 register moves, 12 of them at the loop top of every
        // iteration -- C5 3 585 -> 3 645 Mrays/s, C4 628 -> 638 (-DB2RT_LOOP_WHILE brings the old shape back)
#ifndef B2RT_LOOP_WHILE
        if (active) for (;;) {
#else
        while (active) {
#endif
            // ---- node step ----------------------------------------------------------------------
            if (G.y > 0x00FFFFFFu) {
                if (STATS) { const unsigned am = __activemask(); stNodeLanes++; if (lane == __ffs(am) - 1) stNodeIters++; }
                const unsigned bit = 31u - __clz(G.y);
                G.y &= ~(1u << bit);
                if (G.y > 0x00FFFFFFu) push(sp, G);
                const unsigned slot = (bit - 24u) ^ (octinv4 & 7u);
                const unsigned rel = __popc(G.y & ((1u << slot) - 1u) & 0xFFu);
                const float4* n = p.nodes + (size_t)(G.x + rel) * 5;
                const float4 q0 = ldg4(n), q1 = ldg4(n + 1), q2 = ldg4(n + 2), q3 = ldg4(n + 3), q4 = ldg4(n + 4);

                const unsigned ew = __float_as_uint(q0.w);
                const float ax = __uint_as_float((ew & 0xFFu) << 23) * idirx;
                const float ay = __uint_as_float((ew & 0xFF00u) << 15) * idiry;
                const float az = __uint_as_float((ew & 0xFF0000u) << 7) * idirz;
                const float bx = __fmaf_rn(q0.x - origx, idirx, -ax);
                const float by = __fmaf_rn(q0.y - origy, idiry, -ay);
                const float bz = __fmaf_rn(q0.z - origz, idirz, -az);
                // entry / exit planes by ray direction: no per-child min/max over the two planes of an axis
                const bool nx = idirx < 0.0f, ny = idiry < 0.0f, nz = idirz < 0.0f;     // idir is never zero (2^-80 clamp keeps the sign)
                // WT: the decoded planes t = (1 + q 2^-15) ax + b carry an error of an ulp of |b| (~2^15 cells: 2^-9 of a cell), which the
                // 1/64-cell margin of the quantised boxes covers for ordinary nodes; a node far from the ray's origin relative to its
                // size (|b| >> 2^15 cells is impossible: b is measured from the node's own origin) needs no more.  The slack below
                // covers rcp.approx and the last rounding.
                const float wslack = WT ? (1.0f / 2097152.0f) * fmaxf(fmaxf(fabsf(bx), fabsf(by)), fabsf(bz)) : 0.0f;
                unsigned hitmask = 0;
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    // (Round 1 skipped the upper half for nodes without children there.  With the SAH-optimal collapse 76 % of the
                    //  nodes have more than four children, and at warp level the upper half ran in 19.1 M of 19.1 M steps anyway:
                    //  removing the branch is +1.2 % on C5 and +1.6 % on C4, profiles/r02_explore_c5_opt.log.)
                    const unsigned loX = __float_as_uint(h ? q2.y : q2.x), loY = __float_as_uint(h ? q2.w : q2.z);
                    const unsigned loZ = __float_as_uint(h ? q3.y : q3.x), hiX = __float_as_uint(h ? q3.w : q3.z);
                    const unsigned hiY = __float_as_uint(h ? q4.y : q4.x), hiZ = __float_as_uint(h ? q4.w : q4.z);
                    const unsigned nearX = nx ? hiX : loX, farX = nx ? loX : hiX;
                    const unsigned nearY = ny ? hiY : loY, farY = ny ? loY : hiY;
                    const unsigned nearZ = nz ? hiZ : loZ, farZ = nz ? loZ : hiZ;
                    const unsigned meta4 = __float_as_uint(h ? q1.w : q1.z);
                    const unsigned isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
                    const unsigned innerMask4 = (isInner4 >> 4) * 0xFFu;
                    const unsigned bitIndex4 = (meta4 ^ (octinv4 & innerMask4)) & 0x1F1F1F1Fu;
                    const unsigned childBits4 = (meta4 >> 5) & 0x07070707u;
                    const unsigned m4 = meta4 ^ (octinv4 & innerMask4);         // per byte: bits 0-4 the bit index, bits 5-7 the child's bits
#define B2RT_WIDE_CHILD(J)                                                                                          \
                    {                                                                                               \
                        const float t0x = __fmaf_rn(byteToUnitFloat<J>(nearX, one), ax, bx);                           \
                        const float t0y = __fmaf_rn(byteToUnitFloat<J>(nearY, one), ay, by);                           \
                        const float t0z = __fmaf_rn(byteToUnitFloat<J>(nearZ, one), az, bz);                           \
                        const float t1x = __fmaf_rn(byteToUnitFloat<J>(farX, one) , ax, bx);                            \
                        const float t1y = __fmaf_rn(byteToUnitFloat<J>(farY, one) , ay, by);                            \
                        const float t1z = __fmaf_rn(byteToUnitFloat<J>(farZ, one) , az, bz);                            \
                        const float cmin = fmax3(t0x, t0y, fmaxf(t0z, tmin));                                      \
                        const float cmax = fmin3(t1x, t1y, fminf(t1z, WT ? cullT : hitT));                                      \
                        if (WT ? watertightOverlap(cmin, cmax, wslack) : (cmin <= cmax)) {                          \
                            if (MASK) { const unsigned m = __byte_perm(m4, 0u, 0x4440u + J); hitmask |= __funnelshift_l(0u, m >> 5, m); } \
                            else hitmask |= ((childBits4 >> (8 * J)) & 0xFFu) << ((bitIndex4 >> (8 * J)) & 0xFFu);       \
                        }                                                                                           \
                    }
                    B2RT_WIDE_CHILD(0) B2RT_WIDE_CHILD(1) B2RT_WIDE_CHILD(2) B2RT_WIDE_CHILD(3)
#undef B2RT_WIDE_CHILD
                }
                G = make_uint2(__float_as_uint(q1.x), (hitmask & 0xFF000000u) | (ew >> 24));
                T = make_uint2(__float_as_uint(q1.y), hitmask & 0x00FFFFFFu);
            } else {
                T = G;                                              // a postponed triangle group came off the stack
                G = make_uint2(0u, 0u);
            }

            // ---- triangles of this node ---------------------------------------------------------
            // Postponing (Ylitie et al. 2017, sec. 4.3): if fewer than 1/POSTPONE of the lanes that took this step
            // hold triangles, intersecting now would run the Woop test at a few lanes wide (ncu: 3.6 active threads per
            // triangle-loop instruction without it) -- park the group on the stack and keep traversing instead.
            // The vote is repeated for every triangle of a group (voting once per group: -6 %).  POSTPONE sweep on C5, Mrays/s and
            // node-loop / triangle-loop SIMD efficiency: 1/3 3 257 (66 % / 38 %), 1/4 3 337, 1/5 3 348 (75 % / 28 %), 1/6 3 336, 1/8 3 277,
            // 1/16 3 047 (83 % / 16 %) -- profiles/r02_explore_c5_postpone.log; 1/5 also wins on C4 and C3.  With PARK (turned-away
            // groups wait in a room of their own and the lane traverses on) the balance moves: 1/5 3 523 (84 % / 29 %), 1/4 3 557
            // (83 % / 32 %), 1/3 3 548 (79 % / 39 %), 1/2 3 342 (69 % / 52 %) -- profiles/r02_explore_c5_last_session.log; shipped: 1/4.
            if (PARK && parked && T.y == 0u) { parked--; T = s_park[(kParkDepth == 1 ? 0 : parked) * BLOCK]; }      // no fresh group: offer the parked one
            const int lanesHere = __popc(__activemask());
            if (LEAF == 1) {
                // as trace.cu LEAF == 1: row 0 of the NEXT triangle of the group is requested before the current one is tested
                // (the hit mask names it), Woop test without early-outs -- the same operations in the same order
                int tri = (int)T.x + (T.y ? __ffs(T.y) - 1 : 0);
                float4 v00 = ldg4(p.woop + (size_t)tri * 3);
                // true: the group (what is left of it) was turned away and put aside
                auto turnedAway = [&]() -> bool {
                    if (PARK) {
                        if (__popc(__activemask()) * POSTPONE < lanesHere) {
                            if (parked < kParkDepth) { s_park[(kParkDepth == 1 ? 0 : parked) * BLOCK] = T; parked++; return true; }
                            if (sp < p.pushLimit) { push(sp, T); return true; }
                        }
                        return false;
                    }
                    if (POSTPONE > 0 && __popc(__activemask()) * POSTPONE < lanesHere && sp < p.pushLimit) { push(sp, T); return true; }
                    return false;
                };
                // tests triangle `cur` (row 0 in `c00`), requests row 0 of the group's next triangle into `n00`
                auto triStep = [&](const int cur, const float4& c00, int& nxt, float4& n00) {
                    if (STATS) { const unsigned am = __activemask(); stTriLanes++; if (lane == __ffs(am) - 1) stTriIters++; }
                    T.y &= T.y - 1u;
                    nxt = (int)T.x + (T.y ? __ffs(T.y) - 1 : 0);
                    const float4* w = p.woop + (size_t)cur * 3;
                    n00 = ldg4(p.woop + (size_t)nxt * 3);
                    const float4 v11 = ldg4(w + 1), v22 = ldg4(w + 2);
                    const float Oz = woopPlane(-origx, -origy, -origz, c00);
                    const float invDz = rcp_approx(woopDot(dirx, diry, dirz, c00));
                    const float t = fmulFtz(Oz, invDz);
                    const float Ox = woopPlane(origx, origy, origz, v11);
                    const float Dx = woopDot(dirx, diry, dirz, v11);
                    const float u = ffmaFtz(t, Dx, Ox);
                    const float Oy = woopPlane(origx, origy, origz, v22);
                    const float Dy = woopDot(dirx, diry, dirz, v22);
                    const float v = ffmaFtz(t, Dy, Oy);
                    if (woopAccept<WT>(t, tmin, hitT, u, v, v11, v22, WtRay{ origx, origy, origz, dirx, diry, dirz }, cur, hitIndex, p.triIndex)) {
                        hitT = t; hitIndex = cur;
                        if (WT) cullT = watertightCull(t);
                        if (BARY) { hitU = u; hitV = v; }
                        if (ANY_HIT) { T.y = 0u; G.y = 0u; sp = 0; parked = 0; }
                    }
                };
                if (OPT & 16) {
                    // unrolled twice, the two row-0 register sets swapping roles: no copy of the requested row per triangle (5 of the loop's
                    // 56 instructions); C5 3 659 -> 3 714, C3 3 476 -> 3 506, C4 unchanged (profiles/r02_explore_c5_last_session.log)
                    int tri2; float4 w00;
                    while (T.y != 0u) {
                        if (turnedAway()) break;
                        triStep(tri, v00, tri2, w00);
                        if (T.y == 0u || turnedAway()) break;
                        triStep(tri2, w00, tri, v00);
                    }
                } else
                while (T.y != 0u) {
                    if (turnedAway()) break;
                    int ntri; float4 n00;
                    triStep(tri, v00, ntri, n00);
                    v00 = n00; tri = ntri;
                }
            } else
            while (T.y != 0u) {
                if (POSTPONE > 0 && __popc(__activemask()) * POSTPONE < lanesHere && sp < p.pushLimit) {
                    push(sp, T);
                    break;
                }
                if (STATS) { const unsigned am = __activemask(); stTriLanes++; if (lane == __ffs(am) - 1) stTriIters++; }
                const int k = __ffs(T.y) - 1;
                T.y &= T.y - 1u;
                const int tri = (int)T.x + k;
                const float4* w = p.woop + (size_t)tri * 3;
                const float4 v00 = ldg4(w), v11 = ldg4(w + 1), v22 = ldg4(w + 2);
                // Woop test, same operation order as trace.cu (kepler_dynamic_fetch.cu:336-379)
                const float Oz = woopPlane(-origx, -origy, -origz, v00);
                const float invDz = rcp_approx(woopDot(dirx, diry, dirz, v00));
                const float t = fmulFtz(Oz, invDz);
                if (WT) {
                    const float Ox = woopPlane(origx, origy, origz, v11), Dx = woopDot(dirx, diry, dirz, v11);
                    const float Oy = woopPlane(origx, origy, origz, v22), Dy = woopDot(dirx, diry, dirz, v22);
                    const float u = ffmaFtz(t, Dx, Ox), v = ffmaFtz(t, Dy, Oy);
                    if (woopAccept<true>(t, tmin, hitT, u, v, v11, v22, WtRay{ origx, origy, origz, dirx, diry, dirz }, tri, hitIndex, p.triIndex)) {
                        hitT = t; hitIndex = tri; cullT = watertightCull(t);
                        if (BARY) { hitU = u; hitV = v; }
                        if (ANY_HIT) { T.y = 0u; G.y = 0u; sp = 0; }
                    }
                } else
                if (t > tmin && t < hitT) {
                    const float Ox = woopPlane(origx, origy, origz, v11);
                    const float Dx = woopDot(dirx, diry, dirz, v11);
                    const float u = ffmaFtz(t, Dx, Ox);
                    if (u >= 0.0f) {
                        const float Oy = woopPlane(origx, origy, origz, v22);
                        const float Dy = woopDot(dirx, diry, dirz, v22);
                        const float v = ffmaFtz(t, Dy, Oy);
                        if (v >= 0.0f && faddFtz(u, v) <= 1.0f) {
                            hitT = t; hitIndex = tri;
                            if (BARY) { hitU = u; hitV = v; }
                            if (ANY_HIT) { T.y = 0u; G.y = 0u; sp = 0; }
                        }
                    }
                }
            }

            // ---- next group ---------------------------------------------------------------------
            if (G.y <= 0x00FFFFFFu) {
                if (sp == 0) {
                    if (PARK && parked) { parked--; G = s_park[(kParkDepth == 1 ? 0 : parked) * BLOCK]; }       // all that is left: comes back like a popped group
                    else { active = false; break; }
                } else
                G = pop(sp);
            }
            if (__popc(__activemask()) < threshold) break;          // dynamic fetch: refill the idle lanes
        }
    }

    if (STATS) {
        for (int o = 16; o > 0; o >>= 1) {
            stNodeIters += __shfl_xor_sync(kFullMask, stNodeIters, o);
            stNodeLanes += __shfl_xor_sync(kFullMask, stNodeLanes, o);
            stTriIters += __shfl_xor_sync(kFullMask, stTriIters, o);
            stTriLanes += __shfl_xor_sync(kFullMask, stTriLanes, o);
            stFetches += __shfl_xor_sync(kFullMask, stFetches, o);
        }
        if (lane == 0) {
            atomicAdd(p.stats + 0, stNodeIters); atomicAdd(p.stats + 1, stNodeLanes);
            atomicAdd(p.stats + 2, stTriIters);  atomicAdd(p.stats + 3, stTriLanes);
            atomicAdd(p.stats + 4, stFetches);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// launcher
// ---------------------------------------------------------------------------------------------
template <bool ANY_HIT, bool BARY, bool STATS, int BLOCK, int MIN_BLOCKS, int POSTPONE, int LEAF, int SSLOTS, int CLAIM, int OPT, bool WT = false>
static int launchWideOne(b2rt_ctx* ctx, WideParams& p, cudaStream_t stream)
{
    auto kern = traceWideKernel<ANY_HIT, BARY, STATS, BLOCK, MIN_BLOCKS, POSTPONE, LEAF, SSLOTS, WT, CLAIM, OPT>;
    int blocksPerSm = 0;
    const size_t smem = (size_t)(SSLOTS + ((OPT & 4) ? ((OPT & 8) ? 2 : 1) : 0)) * BLOCK * sizeof(uint2)
                      + (((OPT & 32) && CLAIM > 0) ? (size_t)(BLOCK / 32) * CLAIM * sizeof(unsigned) : 0);
    { const int rc = kernelOccupancy(ctx, kern, BLOCK, smem, &blocksPerSm); if (rc) return rc; }
    if (blocksPerSm < 1) return setError(ctx, B2RT_ERR_CUDA, "wide trace kernel does not fit on an SM");
    if (ctx->cfg.blocks_per_sm > 0 && ctx->cfg.blocks_per_sm < blocksPerSm) blocksPerSm = ctx->cfg.blocks_per_sm;
    long long grid = (long long)ctx->smCount * blocksPerSm;
    const long long needed = ((long long)p.numRays + BLOCK - 1) / BLOCK;
    if (grid > needed) grid = needed;
    if (grid < 1) grid = 1;
    p.warpCounter = ctx->warpCounter + (ctx->launchSlot++ % b2rt_ctx::kLaunchSlots) * 32;
    // stack: SSLOTS entries in shared memory + kWideStack in local memory; sibling groups (at most one per level) must always fit
    p.pushLimit = kWideStack + SSLOTS - ctx->wideMaxDepth - 1;
    if (p.pushLimit < 1) return setError(ctx, B2RT_ERR_ARG, "wide BVH is %d levels deep: too deep for the wide kernel's stack", ctx->wideMaxDepth);
    B2RT_CUDA(ctx, cudaMemsetAsync(p.warpCounter, 0, sizeof(int), stream));
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)grid); lc.blockDim = dim3(BLOCK); lc.dynamicSmemBytes = smem; lc.stream = stream;
    cudaLaunchAttribute attr[1];
    lc.attrs = attr; lc.numAttrs = 0;
    if ((ctx->cfg.tune & B2RT_TUNE_L2_WINDOW) && ctx->l2SetAside >= ctx->wideBytes) {
        // the whole wide form (nodes, Woop rows, ids: one allocation) stays in the set-aside part of L2; rays and records stream past it
        attr[0].id = cudaLaunchAttributeAccessPolicyWindow;
        attr[0].val.accessPolicyWindow.base_ptr = ctx->wideNodes;
        attr[0].val.accessPolicyWindow.num_bytes = ctx->wideBytes;
        attr[0].val.accessPolicyWindow.hitRatio = 1.0f;
        attr[0].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr[0].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        lc.numAttrs = 1;
    }
    B2RT_CUDA(ctx, cudaLaunchKernelEx(&lc, kern, p));
    B2RT_CHECK_LAUNCH(ctx, "traceWideKernel");
    return B2RT_OK;
}

template <int BLOCK, int MIN_BLOCKS, int POSTPONE, int LEAF = 0, int SSLOTS = 0, int CLAIM = 0, int OPT = 0>
static int launchWideFlags(b2rt_ctx* ctx, WideParams& p, unsigned flags, cudaStream_t stream)
{
    const bool any = (flags & B2RT_TRACE_ANY_HIT) != 0;
    const bool bary = (flags & B2RT_TRACE_WRITE_BARY) != 0;
    if (flags & B2RT_TRACE_WATERTIGHT)
        return any ? launchWideOne<true, true, false, BLOCK, MIN_BLOCKS, POSTPONE, LEAF, SSLOTS, CLAIM, OPT, true>(ctx, p, stream)
                   : launchWideOne<false, true, false, BLOCK, MIN_BLOCKS, POSTPONE, LEAF, SSLOTS, CLAIM, OPT, true>(ctx, p, stream);
    if (flags & B2RT_TRACE_STATS)
        return any ? launchWideOne<true, true, true, BLOCK, MIN_BLOCKS, POSTPONE, LEAF, SSLOTS, CLAIM, OPT>(ctx, p, stream)
                   : launchWideOne<false, true, true, BLOCK, MIN_BLOCKS, POSTPONE, LEAF, SSLOTS, CLAIM, OPT>(ctx, p, stream);
    if (any) return bary ? launchWideOne<true, true, false, BLOCK, MIN_BLOCKS, POSTPONE, LEAF, SSLOTS, CLAIM, OPT>(ctx, p, stream)
                         : launchWideOne<true, false, false, BLOCK, MIN_BLOCKS, POSTPONE, LEAF, SSLOTS, CLAIM, OPT>(ctx, p, stream);
    return bary ? launchWideOne<false, true, false, BLOCK, MIN_BLOCKS, POSTPONE, LEAF, SSLOTS, CLAIM, OPT>(ctx, p, stream)
                : launchWideOne<false, false, false, BLOCK, MIN_BLOCKS, POSTPONE, LEAF, SSLOTS, CLAIM, OPT>(ctx, p, stream);
}

int launchTraceWide(b2rt_ctx* ctx, const void* rays, void* results, int64_t numRays, unsigned flags, int variant,
                    cudaStream_t stream)
{
    int rc = ensureWide(ctx);
    if (rc) return rc;
    if ((ctx->cfg.tune & B2RT_TUNE_L2_WINDOW) && ctx->l2SetAside == 0) {
        // ask once for as much persisting L2 as the device gives (device-wide limit; other kernels simply do not use it)
        cudaDeviceProp prop;
        B2RT_CUDA(ctx, cudaGetDeviceProperties(&prop, ctx->device));
        B2RT_CUDA(ctx, cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)prop.persistingL2CacheMaxSize));
        ctx->l2SetAside = (size_t)prop.persistingL2CacheMaxSize;
    }
    WideParams p;
    p.rays = static_cast<const float4*>(rays);
    p.results = static_cast<int4*>(results);
    p.nodes = ctx->wideNodes;
    p.woop = ctx->wideWoop;
    p.triIndex = ctx->wideTriIndex;
    p.numRays = (int)numRays;
    // refill below 24 live lanes (the binary kernel and the reference use 20): a wide step is ~3x as many instructions as a
    // binary one, so an idle lane costs more; measured 12 / 16 / 20 / 24 / 28 on C5: 3 119 / 3 205 / 3 203 / 3 241 / 3 175,
    // C3: 3 086 (20) -> 3 153 (24)  (profiles/r01_explore_c{3,5}_wide_threshold.log)
    p.fetchThreshold = ctx->cfg.fetch_threshold < 0 ? ((variant == 114 || variant == 124 || variant == 142 || variant == 145 || variant == 147 || variant == 148 || variant == 160 || variant == 162 || variant == 170 || variant == 180 || variant == 181 || (variant >= 190 && variant <= 194)) ? 26 : 24) : ctx->cfg.fetch_threshold;
    p.warpCounter = ctx->warpCounter;
    p.stats = ctx->statsDev;
    p.one = 0x3F800000u;
    p.streamRays = (ctx->cfg.tune & B2RT_TUNE_NO_STREAM) ? 0 : 1;
    p.pushLimit = 0;
    if (flags & B2RT_TRACE_STATS) B2RT_CUDA(ctx, cudaMemsetAsync(ctx->statsDev, 0, 16 * sizeof(unsigned long long), stream));
    switch (variant) {
    default:
    case 108: return launchWideFlags<256, 4, 4, 1, 8, 0, 22>(ctx, p, flags, stream);   // <= 64 regs, 32 warps/SM; look-ahead triangle loop unrolled twice, 8 stack
                                                                                       // entries in shared memory; triangle groups turned away below 1/4 of the lanes
                                                                                       // wait in a one-entry room of their own (PARK), hit mask through PRMT (MASK)
    case 114: return launchWideFlags<256, 4, 4, 1, 8, 128, 54>(ctx, p, flags, stream); // 108 with 128-ray block claims, each block classified at once (SCAN):
                                                                                       // batches above 12 Mi rays
#ifdef B2RT_EXPLORE
    // shapes measured and not shipped; build with -DB2RT_EXPLORE to get them back
    case 160: return launchWideFlags<256, 4, 4, 1, 8, 128, 14>(ctx, p, flags, stream);   // 114 with two waiting-room entries: C5 3 585 -> 3 480, C3 3 434 -> 3 344
    case 161: return launchWideFlags<256, 4, 4, 1, 8, 0, 14>(ctx, p, flags, stream);     // 108 with two
    case 162: return launchWideFlags<256, 4, 3, 1, 8, 128, 14>(ctx, p, flags, stream);   //   ... threshold 1/3
    case 190: return launchWideFlags<256, 4, 4, 1, 6, 128, 54>(ctx, p, flags, stream);   // 114 with 6 shared stack slots: C5 3 802 -> 3 801
    case 191: return launchWideFlags<256, 4, 4, 1, 10, 128, 54>(ctx, p, flags, stream);  //   ... 10: 3 788
    case 192: return launchWideFlags<256, 4, 4, 1, 12, 128, 54>(ctx, p, flags, stream);  //   ... 12: 3 787
    case 193: return launchWideFlags<256, 4, 3, 1, 8, 128, 54>(ctx, p, flags, stream);   // 114 with the threshold at 1/3: 3 796
    case 194: return launchWideFlags<256, 4, 5, 1, 8, 128, 54>(ctx, p, flags, stream);   //   ... 1/5: 3 768
    case 180: return launchWideFlags<256, 4, 4, 1, 8, 128, 22>(ctx, p, flags, stream);   // 114 without SCAN: C5 3 785 -> 3 719
    case 181: return launchWideFlags<256, 4, 4, 1, 8, 256, 54>(ctx, p, flags, stream);   // 114 with 256-ray blocks: 3 728
    case 182: return launchWideFlags<256, 4, 4, 1, 8, 32, 54>(ctx, p, flags, stream);    // 32-ray blocks + SCAN on frame-sized batches: C4 662 -> 636, C3 3 484 -> 3 293
    case 183: return launchWideFlags<256, 4, 4, 1, 8, 64, 54>(ctx, p, flags, stream);    // 64-ray blocks: C4 609, C3 3 260; 8 Mi-ray shard of C5 3 419 -> 3 437
    case 170: return launchWideFlags<256, 4, 4, 1, 8, 128, 6>(ctx, p, flags, stream);    // 114 with the triangle loop not unrolled: C5 3 714 -> 3 659, C3 3 506 -> 3 476
    case 171: return launchWideFlags<256, 4, 4, 1, 8, 0, 6>(ctx, p, flags, stream);      // 108 ...
    case 118: return launchWideFlags<256, 4, 5, 1, 8>(ctx, p, flags, stream);            // variant 108 until the end of round 2: groups pushed on the stack, threshold 1/5
    case 124: return launchWideFlags<256, 4, 5, 1, 8, 128>(ctx, p, flags, stream);       // variant 114 until then
    case 142: return launchWideFlags<256, 4, 5, 1, 8, 128, 2>(ctx, p, flags, stream);    // 124 + MASK
    case 145: return launchWideFlags<256, 4, 5, 1, 8, 128, 6>(ctx, p, flags, stream);    // 124 + MASK + PARK, threshold 1/5
    case 147: return launchWideFlags<256, 4, 3, 1, 8, 128, 6>(ctx, p, flags, stream);    //   ... 1/3
    case 148: return launchWideFlags<256, 4, 2, 1, 8, 128, 6>(ctx, p, flags, stream);    //   ... 1/2
    case 150: return launchWideFlags<256, 4, 3, 1, 8, 0, 6>(ctx, p, flags, stream);      // 108 with the threshold at 1/3
    case 100: return launchWideFlags<256, 3, 0>(ctx, p, flags, stream);   // <= 80 regs, 24 warps/SM, no postponing
    case 104: return launchWideFlags<256, 4, 3>(ctx, p, flags, stream);   // postpone below 1/3, reference-shaped triangle loop
    case 106: return launchWideFlags<256, 4, 3, 1>(ctx, p, flags, stream);     // 104 with the look-ahead triangle loop
    case 107: return launchWideFlags<256, 4, 3, 1, 8>(ctx, p, flags, stream);  // round 1's default (postpone below 1/3)
    case 109: return launchWideFlags<256, 4, 5, 1, 12>(ctx, p, flags, stream); // 12 stack entries in shared memory
    case 111: return launchWideFlags<256, 4, 0, 1, 8>(ctx, p, flags, stream);  // 108 without postponing
    case 113: return launchWideFlags<512, 2, 5, 1, 8>(ctx, p, flags, stream);  // 108 in two 512-thread CTAs
    case 115: return launchWideFlags<256, 5, 5, 1, 8>(ctx, p, flags, stream);  // <= 48 regs (spills), 40 warps/SM: -7 %
    case 132: return launchWideFlags<256, 4, 5, 1, 8, 64>(ctx, p, flags, stream);   // 64-ray claims
    case 134: return launchWideFlags<256, 4, 5, 1, 8, 256>(ctx, p, flags, stream);  // 256-ray claims
#endif
    }
}

}  // namespace b2rt
