// trace.cu -- the hot path: batched closest-hit / any-hit traversal of a Compact2 BVH.
//
// Replaces rt/kernels/kepler_dynamic_fetch.cu:66-411 (`trace`) and its launcher :432-474.
// Written from scratch for sm_100a; what it shares with the reference is the contract
// (record layouts, child encoding, leaf terminator, result semantics), not the code:
//
//  * persistent warps pull rays through one global counter; the refill is aggregated per warp
//    with __ballot_sync/__popc/__shfl_sync (no shared-memory mailbox, no volatile), and every
//    lane stays in the loop until the whole warp is out of work, so all votes use a full mask;
//  * while-while traversal with one postponed leaf (speculative traversal) and a dynamic
//    refill when fewer than `fetchThreshold` lanes are still traversing;
//  * nodes (64 B) and Woop triangles (48 B) are read with 128-bit loads; the first `topF4/4`
//    nodes of the re-laid node array (breadth-first top of the tree) are staged in shared
//    memory once per CTA and read from there;
//  * the traversal stack keeps its top entry in a register, the next STACK_SMEM entries in
//    shared memory (bank = lane, conflict free) and spills the rest to local memory;
//  * slab test with 3-input FMNMX3 min/max, Woop test with explicit FMA order and
//    rcp.approx (the reference is built with --use_fast_math, Makefile:120);
//  * a result is stored exactly once per ray (the reference re-stores partial results at every
//    dynamic-fetch exit, kepler_dynamic_fetch.cu:400-408), 16 B with barycentrics or 8 B without.
#include "common.cuh"

namespace b2rt {

#ifdef B2RT_EXPLORE
constexpr bool kAlwaysStream = false;       // B2RT_TUNE_NO_STREAM is an A/B switch in -DB2RT_EXPLORE builds only (trace_wide.cu)
#else
constexpr bool kAlwaysStream = true;
#endif

struct TraceParams {
    const float4* __restrict__ rays;
    int4* __restrict__ results;
    const float4* __restrict__ nodes;
    const float4* __restrict__ woop;
    const int* __restrict__ triIndex;
    int numRays;
    int topF4;              // float4 elements of the node array staged in shared memory
    int fetchThreshold;
    int dummyLeaf;          // ~row of the terminator row behind the caller's Woop rows: a leaf without triangles (SHARE)
    int* shareList;         // SHARE: kShareList ray indices per CTA
    int streamRays;         // 1: rays are read and records written with the streaming (evict-first) policy: they pass once, the BVH stays in L2
    int* warpCounter;
    unsigned long long* stats;
};

constexpr int kLocalStack = 64;     // SplitBVHBuilder MaxDepth = 64 bounds the stack (SplitBVHBuilder.hh:43)
constexpr unsigned kFull = 0xffffffffu;

// LEAF: shape of the leaf loop.  0 = three rows per triangle fetched together, terminator tested on the fetched row 0, Woop
// test with early-outs (the reference's shape, kepler_dynamic_fetch.cu:326-379).  1 = row 0 of the next triangle requested
// ahead, test without early-outs (below).  Measured (profiles/r01_explore_c{2,3,5}_leaf_ahead.log): LEAF 1 wins on frame-sized
// batches, C2 4 583 -> 4 753 and C3 2 753 -> 2 890 Mrays/s, and loses on the 64 Mi-ray incoherent batch, C5 3 001 -> 2 797.
// SHARE (compiled into the ANY_HIT kernels): subtree sharing in the tail of a launch.  Once the work counter has run dry a
// launch is a set of dependent chains -- every remaining ray needs its remaining node visits one after the other, ~380 ns
// each -- while most lanes and warps of the SM have nothing left to do (C3 AO batches: pool empty at 285 us, last warp at
// 440 us).  An any-hit ray that misses (96 % of C3's) has to visit EVERY entry of its stack, so the entries are real,
// independent work: lanes that are still traversing hand the LOWEST entries of their stacks -- whole subtrees -- to a
// queue in shared memory; idle lanes of any warp of the CTA take them, re-read the ray, traverse the subtree and merge a
// hit into the ray's result record with a 64-bit atomicMin on (t, Woop row).  After the CTA's last warp has left the loop
// the CTA rewrites the records of its shared rays into the final form (triangle id, barycentrics).  Nothing waits on
// anything but the queue's try-lock.  Closest-hit kernels do not share: front to back, the old stack entries are far
// subtrees the owner culls once it has its hit, and handing them over bought nothing.
struct ShareArea {
    int lock;               // 0 free, 1 taken (try-lock; idle warps retry, busy warps just skip the visit)
    int head, tail;         // queue positions, free-running; changed under the lock
    int idleWarps;          // warps of the CTA without a ray in flight (includes warps that have left)
    int listCount;          // shared rays of this CTA (entries of its slice of TraceParams::shareList)
    int pad[3];
};
constexpr int kShareSteps = 4;      // tail of the launch: node steps between two visits of the queue
constexpr int kShareRounds = 4;     // stack entries a lane may hand over per visit
constexpr int kShareQueue = 128;    // queue entries (int4: ray, stack entry, hit distance); 2 KB keeps the L1 carve-out where it was
constexpr int kShareList = 1024;    // shared rays per CTA (global memory, one slice per CTA and launch slot)

template <bool ANY_HIT, bool BARY, bool STATS, int STACK_SMEM, int BLOCK, int MIN_BLOCKS, int LEAF, bool WT>
__global__ void __launch_bounds__(BLOCK, MIN_BLOCKS) traceKernel(const TraceParams p)
{
    extern __shared__ float4 s_dyn[];
    // Shared memory: [stack: STACK_SMEM x BLOCK ints][top nodes, structure-of-arrays].
    // The top `topNodes` nodes of the re-laid array are staged as four planes -- plane 0/1/2 hold the
    // node's first three float4s (16 B per node), plane 3 its two child indices (8 B per node) -- so
    // lanes that read DIFFERENT nodes in one LDS.128 hit banks (node mod 8) instead of all starting
    // at bank 0 or 16 as an array-of-64-byte-structs would (ncu: 8.6 M excessive shared wavefronts
    // per launch in the AoS version, profiles/r01_ncu_trace_c2.md).
    int* s_stack = reinterpret_cast<int*>(s_dyn) + threadIdx.x;              // slot k at s_stack[k * BLOCK]
    const int topF4 = p.topF4;
    const int topNodes = topF4 >> 2;
    float4* s_plane = s_dyn + (STACK_SMEM * BLOCK) / 4;
    {
        int2* s_child = reinterpret_cast<int2*>(s_plane + 3 * topNodes);
        for (int i = threadIdx.x; i < topF4; i += BLOCK) {
            const float4 v = ldg4(p.nodes + i);
            const int node = i >> 2, k = i & 3;
            if (k < 3) s_plane[k * topNodes + node] = v;
            else s_child[node] = make_int2(__float_as_int(v.x), __float_as_int(v.y));
        }
    }
    // SHARE: [ShareArea][queue] behind the planes
    constexpr bool SHARE = ANY_HIT;
    constexpr int kQueue = kShareQueue, kListCap = kShareList, kWarps = BLOCK / 32;
    ShareArea* s_share = reinterpret_cast<ShareArea*>(reinterpret_cast<char*>(s_plane) + (size_t)topNodes * 56);
    int4* s_queue = reinterpret_cast<int4*>(s_share + 1);
    int* s_list = p.shareList + (size_t)blockIdx.x * kListCap;
    if (SHARE && threadIdx.x == 0) { s_share->lock = 0; s_share->head = 0; s_share->tail = 0; s_share->idleWarps = 0; s_share->listCount = 0; }
    __syncthreads();
    unsigned sTop = (unsigned)__cvta_generic_to_shared(s_plane);              // 32-bit shared address of plane 0
    asm volatile("mov.u32 %0, %0;" : "+r"(sTop));      // opaque: keep it in a register instead of re-deriving it per node
    const unsigned planeBytes = (unsigned)topNodes * 16u;

    const int lane = threadIdx.x & 31;
    int lstack[kLocalStack];            // spill area, touched only when a stack outgrows STACK_SMEM

    // Per-ray state.
    float origx = 0, origy = 0, origz = 0, dirx = 0, diry = 0, dirz = 0;
    float idirx = 0, idiry = 0, idirz = 0, oodx = 0, oody = 0, oodz = 0, tmin = 0, hitT = 0;
    float hitU = 0, hitV = 0;
    float cullT = 0, slack = 0;         // WT: hitT widened for the box tests, absolute slack of the slab distances (common.cuh)
    int hitIndex = -1, rayidx = -1;
    int nodeAddr = kSentinel;           // >= 0 inner node, < 0 leaf, sentinel = finished
    int leafAddr = 0;                   // postponed leaf (< 0) or none (>= 0)
    int tos = kSentinel;                // top of stack, in a register
    int sp = 0;                         // number of entries held in memory
    bool poolEmpty = false;             // warp-uniform: the work counter ran past numRays

    // subtree sharing (SHARE)
    bool viaAtomic = false;             // this lane's ray is shared (or the lane works on a handed-over subtree): merge by atomicMin
    int giveSlot = 2;                   // lowest stack slot not yet handed over (slots 0 and 1 hold the sentinel)
    bool countedIdle = false;           // warp-uniform: this warp is counted in ShareArea::idleWarps

    unsigned long long stNodeIters = 0, stNodeLanes = 0, stTriIters = 0, stTriLanes = 0, stFetches = 0;
    unsigned long long stStart = 0, stPoolEmpty = 0, stShares = 0, stLastFetch = 0, stItersAtFetch = 0;
    if (STATS) stStart = globalTimerNs();

    auto warpSum = [&](unsigned long long v) { for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o); return v; };
    auto stackStore = [&](int slot, int v) {

        if (slot < STACK_SMEM) s_stack[slot * BLOCK] = v; else lstack[slot - STACK_SMEM] = v;
    };
    auto stackLoad = [&](int slot) -> int {
        return (slot < STACK_SMEM) ? s_stack[slot * BLOCK] : lstack[slot - STACK_SMEM];
    };
    auto push = [&](int v) { stackStore(sp, tos); sp++; tos = v; };
    auto pop = [&]() -> int { int r = tos; sp--; tos = stackLoad(sp); return r; };

    while (true) {
        __syncwarp();
        // ---- store finished rays, refill -------------------------------------------------
        const bool terminated = (nodeAddr == kSentinel);
        if (terminated && rayidx >= 0) {
            if (SHARE && viaAtomic) {
                // shared ray: (t, Woop row) competes with what the other lanes working on this ray found
                if (hitIndex != -1)
                    atomicMin(reinterpret_cast<unsigned long long*>(p.results + rayidx),
                              ((unsigned long long)__float_as_uint(hitT) << 32) | (unsigned)hitIndex);
            } else {
                const int id = (hitIndex == -1) ? -1 : __ldg(p.triIndex + hitIndex);
                if (BARY) {
                    const int4 r = make_int4(id, __float_as_int(hitT), __float_as_int(hitU), __float_as_int(hitV));
                    if (kAlwaysStream || p.streamRays) __stcs(p.results + rayidx, r); else p.results[rayidx] = r;
                } else {
                    int2* dst = reinterpret_cast<int2*>(p.results + rayidx);      // 8 of the record's 16 bytes (CudaTracerKernels.hh:197)
                    const int2 r = make_int2(id, __float_as_int(hitT));
                    if (kAlwaysStream || p.streamRays) __stcs(dst, r); else *dst = r;
                }
            }
            rayidx = -1;
        }
        const unsigned maskT = __ballot_sync(kFull, terminated);
        // A warp of this CTA without work means the pool is empty: start visiting the queue.  poolEmpty has to stay
        // warp-uniform (it guards full-mask shuffles) and the lanes need not have reconverged after the stores above, so
        // they could see different values of the counter: lane 0 reads, everybody takes its answer.
        if (SHARE && !poolEmpty) {
            int idle = 0;
            if (lane == 0) idle = *reinterpret_cast<volatile int*>(&s_share->idleWarps);
            if (__shfl_sync(kFull, idle, 0) > 0) poolEmpty = true;
        }
        if (maskT != 0 && !poolEmpty) {
            // unsigned arithmetic: the counter runs past numRays by up to one fetch per warp, which must not wrap for batches near 2^31 rays
            const unsigned numT = __popc(maskT);
            const int leader = __ffs(maskT) - 1;
            unsigned base = 0;
            if (lane == leader) base = (unsigned)atomicAdd(p.warpCounter, (int)numT);
            base = __shfl_sync(kFull, base, leader);
            if (STATS && lane == leader) stFetches++;
            if (STATS && base < (unsigned)p.numRays) { stLastFetch = globalTimerNs(); stItersAtFetch = warpSum(stNodeIters) + (warpSum(stTriIters) << 32); }
            poolEmpty = (base + numT >= (unsigned)p.numRays);
            if (STATS && poolEmpty) stPoolEmpty = globalTimerNs();
            if (terminated) {
                const unsigned idx = base + __popc(maskT & ((1u << lane) - 1u));
                if (idx < (unsigned)p.numRays) {
                    rayidx = (int)idx;
                    float4 o, d;
                    if (kAlwaysStream || p.streamRays) { o = __ldcs(p.rays + 2 * (size_t)idx); d = __ldcs(p.rays + 2 * (size_t)idx + 1); }
                    else { o = ldg4(p.rays + 2 * (size_t)idx); d = ldg4(p.rays + 2 * (size_t)idx + 1); }
                    origx = o.x; origy = o.y; origz = o.z; tmin = o.w;
                    dirx = d.x; diry = d.y; dirz = d.z; hitT = d.w;
                    if (WT) cullT = hitT;
                    const float ooeps = 8.271806125530277e-25f;    // 2^-80, avoids division by zero
                    idirx = rcp_approx(fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x));
                    idiry = rcp_approx(fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y));
                    idirz = rcp_approx(fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z));
                    oodx = origx * idirx; oody = origy * idiry; oodz = origz * idirz;
                    if (WT) slack = watertightSlack(oodx, oody, oodz);
                    // stack: memory slot 0 and the register both hold the sentinel
                    stackStore(0, kSentinel);
                    sp = 1; tos = kSentinel;
                    leafAddr = 0; nodeAddr = 0; hitIndex = -1;
                    if (SHARE) { giveSlot = 2; viaAtomic = false; }
                }
            }
        }
        if (SHARE && poolEmpty) {
            bool leave = false;
            for (;;) {
                const bool busy = (nodeAddr != kSentinel);
                const unsigned busyMask = __ballot_sync(kFull, busy);
                const unsigned idleMask = ~busyMask;
                // hand over only from rays whose t is a positive float (its bits then order like the value)
                if (giveSlot > sp) giveSlot = max(sp, 2);           // the stack has shrunk below entries handed over earlier
                const bool canGive = busy && sp > giveSlot && tmin >= 0.0f;
                const unsigned giveMask = __ballot_sync(kFull, canGive);
                volatile ShareArea* sh = s_share;
                // hints, read by lane 0 for the whole warp (every decision below has to be warp-uniform); the exact
                // values are read under the lock
                int qLen = 0, idleW = 0;
                if (lane == 0) { qLen = sh->tail - sh->head; idleW = sh->idleWarps; }
                qLen = __shfl_sync(kFull, qLen, 0); idleW = __shfl_sync(kFull, idleW, 0);
                const bool wantPop = idleMask != 0 && qLen > 0;
                const bool wantPush = giveMask != 0 && qLen < kQueue / 2 && (idleMask != 0 || idleW > 0);
                if (wantPop || wantPush) {
                    int got = 0;
                    if (lane == 0) got = (atomicCAS(&s_share->lock, 0, 1) == 0);
                    got = __shfl_sync(kFull, got, 0);
                    if (got) {
                        __threadfence_block();
                        int h = 0, t = 0, listCount = 0;
                        if (lane == 0) { h = sh->head; t = sh->tail; listCount = sh->listCount; }
                        h = __shfl_sync(kFull, h, 0); t = __shfl_sync(kFull, t, 0); listCount = __shfl_sync(kFull, listCount, 0);
                        const unsigned below = (1u << lane) - 1u;
                        if (wantPush) {
                            // Up to kShareRounds entries per lane and visit, lowest first: with most of the SM idle the
                            // subtrees should fan out as fast as the queue takes them.
                            bool anyFresh = false;
                            for (int round = 0; round < kShareRounds; round++) {
                                const bool give = busy && sp > giveSlot && tmin >= 0.0f;
                                // lanes whose ray is not shared yet also need a slot in the CTA's list of shared rays
                                const bool needList = give && !viaAtomic;
                                const unsigned needMask = __ballot_sync(kFull, needList);
                                const bool listOk = !needList || listCount + __popc(needMask & below) < kListCap;
                                const unsigned okMask = __ballot_sync(kFull, give && listOk);
                                const int space = kQueue - (t - h);
                                if (okMask == 0 || space <= 0) break;
                                const int rank = __popc(okMask & below);
                                bool fresh = false;
                                if (give && listOk && rank < space) {
                                    if (!viaAtomic) {
                                        s_list[listCount + __popc(needMask & below)] = rayidx;
                                        *reinterpret_cast<unsigned long long*>(p.results + rayidx) = ~0ull;     // "nothing found yet"
                                        viaAtomic = true; fresh = true;
                                    }
                                    const int entry = stackLoad(giveSlot);
                                    stackStore(giveSlot, p.dummyLeaf);
                                    giveSlot++;
                                    s_queue[(unsigned)(t + rank) % kQueue] = make_int4(rayidx, entry, __float_as_int(hitT), 0);
                                    if (STATS) stShares++;
                                }
                                // (the lanes that give are the first ones in lane order, so the new list slots are contiguous)
                                const unsigned gaveFresh = __ballot_sync(kFull, fresh);
                                listCount += __popc(gaveFresh);
                                anyFresh |= gaveFresh != 0;
                                t += min(__popc(okMask), space);
                            }
                            if (anyFresh) __threadfence();          // the cleared records are in L2 before anybody can take the entries
                            __syncwarp();
                        }
                        const int nPop = min(__popc(idleMask), t - h);
                        const int prank = __popc(idleMask & below);
                        int4 item = make_int4(-1, 0, 0, 0);
                        if (!busy && prank < nPop) item = s_queue[(unsigned)(h + prank) % kQueue];
                        __syncwarp();
                        if (lane == 0) {
                            sh->head = h + nPop; sh->tail = t; sh->listCount = listCount;
                            if (nPop > 0 && countedIdle) atomicSub(&s_share->idleWarps, 1);
                            __threadfence_block();
                            atomicExch(&s_share->lock, 0);
                        }
                        if (nPop > 0) countedIdle = false;
                        if (item.x >= 0) {
                            // a handed-over subtree: the ray again, the hit distance of the moment it was handed over
                            rayidx = item.x;
                            const float4 o = ldg4(p.rays + 2 * (size_t)rayidx);
                            const float4 d = ldg4(p.rays + 2 * (size_t)rayidx + 1);
                            origx = o.x; origy = o.y; origz = o.z; tmin = o.w;
                            dirx = d.x; diry = d.y; dirz = d.z; hitT = __int_as_float(item.z);
                            if (WT) cullT = hitT;
                            const float ooeps = 8.271806125530277e-25f;
                            idirx = rcp_approx(fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x));
                            idiry = rcp_approx(fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y));
                            idirz = rcp_approx(fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z));
                            oodx = origx * idirx; oody = origy * idiry; oodz = origz * idirz;
                            if (WT) slack = watertightSlack(oodx, oody, oodz);
                            hitIndex = -1; viaAtomic = true; giveSlot = 2;
                            stackStore(0, kSentinel);
                            sp = 1; tos = kSentinel;
                            // an inner node starts the node loop; a leaf goes straight to the postponed slot, with the empty
                            // leaf behind it so that the loops below run their usual course
                            if (item.y >= 0) { nodeAddr = item.y; leafAddr = 0; } else { leafAddr = item.y; nodeAddr = p.dummyLeaf; }
                        }
                    }
                }
                if (__ballot_sync(kFull, nodeAddr != kSentinel) != 0) break;        // somebody has work: traverse
                // the whole warp is idle: say so, then wait for entries or for everybody else to run dry as well
                if (!countedIdle) {
                    if (lane == 0) atomicAdd(&s_share->idleWarps, 1);
                    countedIdle = true;
                }
                {
                    int done = 0;
                    if (lane == 0) done = (sh->tail - sh->head <= 0 && sh->idleWarps >= kWarps);
                    if (__shfl_sync(kFull, done, 0)) { leave = true; break; }
                }
                __nanosleep(200);
            }
            if (leave) break;
        }
        if (!SHARE || !poolEmpty) { if (__ballot_sync(kFull, nodeAddr != kSentinel) == 0) break; }     // warp out of work

        // after the pool is empty: without SHARE never leave the loop early; with SHARE come back up after every pass
        const int threshold = poolEmpty ? (SHARE ? 33 : 0) : p.fetchThreshold;

        // ---- traversal --------------------------------------------------------------------
        while (nodeAddr != kSentinel) {
            // inner nodes, until every participating lane holds a postponed leaf
            int tailSteps = kShareSteps;        // SHARE, pool empty: go back up for sharing every few node steps
            while ((unsigned)nodeAddr < (unsigned)kSentinel) {
                if (STATS) { const unsigned am = __activemask(); stNodeLanes++; if (lane == __ffs(am) - 1) stNodeIters++; }
                float4 n0xy, n1xy, nz;
                int c0, c1;
                if (nodeAddr < topF4) {
                    const unsigned a = sTop + ((unsigned)nodeAddr << 2);      // node index * 16 bytes
                    n0xy = lds128(a); n1xy = lds128(a + planeBytes); nz = lds128(a + 2u * planeBytes);
                    const int2 ch = lds64i(sTop + 3u * planeBytes + ((unsigned)nodeAddr << 1));
                    c0 = ch.x; c1 = ch.y;
                } else {
                    const float4* np = p.nodes + nodeAddr;
                    n0xy = ldg4(np); n1xy = ldg4(np + 1); nz = ldg4(np + 2);
                    const int2 ch = __ldg(reinterpret_cast<const int2*>(np + 3));
                    c0 = ch.x; c1 = ch.y;
                }

                const float c0lox = __fmaf_rn(n0xy.x, idirx, -oodx), c0hix = __fmaf_rn(n0xy.y, idirx, -oodx);
                const float c0loy = __fmaf_rn(n0xy.z, idiry, -oody), c0hiy = __fmaf_rn(n0xy.w, idiry, -oody);
                const float c0loz = __fmaf_rn(nz.x, idirz, -oodz),   c0hiz = __fmaf_rn(nz.y, idirz, -oodz);
                const float c1loz = __fmaf_rn(nz.z, idirz, -oodz),   c1hiz = __fmaf_rn(nz.w, idirz, -oodz);
                const float c1lox = __fmaf_rn(n1xy.x, idirx, -oodx), c1hix = __fmaf_rn(n1xy.y, idirx, -oodx);
                const float c1loy = __fmaf_rn(n1xy.z, idiry, -oody), c1hiy = __fmaf_rn(n1xy.w, idiry, -oody);
                const float c0min = fmax3(fminf(c0lox, c0hix), fminf(c0loy, c0hiy), fmaxf(fminf(c0loz, c0hiz), tmin));
                const float c0max = fmin3(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy), fminf(fmaxf(c0loz, c0hiz), WT ? cullT : hitT));
                const float c1min = fmax3(fminf(c1lox, c1hix), fminf(c1loy, c1hiy), fmaxf(fminf(c1loz, c1hiz), tmin));
                const float c1max = fmin3(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy), fminf(fmaxf(c1loz, c1hiz), WT ? cullT : hitT));

                const bool trav0 = WT ? watertightOverlap(c0min, c0max, slack) : (c0max >= c0min);
                const bool trav1 = WT ? watertightOverlap(c1min, c1max, slack) : (c1max >= c1min);
                if (!trav0 && !trav1) {
                    nodeAddr = pop();
                } else {
                    nodeAddr = trav0 ? c0 : c1;
                    if (trav0 && trav1) {
                        if (c1min < c0min) { const int t = nodeAddr; nodeAddr = c1; c1 = t; }
                        push(c1);
                    }
                }
                // first leaf: postpone it and keep traversing (measured: folding the two pops into one site, with the popped
                // leaf postponed after the loop, costs 4 % on the diffuse batch -- the lane stops speculating earlier)
                if (nodeAddr < 0 && leafAddr >= 0) {
                    leafAddr = nodeAddr;
                    nodeAddr = pop();
                }
                // every lane still here holds a leaf -> go intersect
                if (!__any_sync(__activemask(), leafAddr >= 0)) break;
                if (SHARE && poolEmpty && --tailSteps == 0) break;
            }

            // postponed leaves
            while (leafAddr < 0) {
                if (LEAF >= 1) {
                    constexpr bool UNROLL = (LEAF == 2);
                    // Row 0 of the NEXT triangle (or the terminator) is requested before the current triangle is tested: the
                    // terminator test and the distance test of the next step need nothing else, so they no longer wait a whole
                    // fetch behind the current test (ncu source page of the LEAF == 0 kernel: 14 % of all stall samples sit on the
                    // terminator compares), and rows 1 and 2 behind the terminator are never read (two fetches fewer per leaf).
                    // The Woop test runs without early-outs: with ~10 lanes in a triangle step one of them nearly always passes
                    // the distance test, so the u and v stages run anyway -- and with conditional uses ptxas sinks the fetches
                    // of rows 1 and 2 behind the distance test (three dependent fetches for a hit; measured slower).  Same
                    // operations in the same order as LEAF == 0: the same bits.
                    int triAddr = ~leafAddr;
                    float4 v00 = ldg4(p.woop + triAddr);
                    // tests the triangle at `addr` (row 0 in c00) and requests row 0 of the next one into n00; true: an any-hit ray is done
                    auto leafStep = [&](const int addr, const float4& c00, float4& n00) -> bool {
                        n00 = ldg4(p.woop + addr + 3);
                        const float4 v11 = ldg4(p.woop + addr + 1), v22 = ldg4(p.woop + addr + 2);
                        if (STATS) { const unsigned am = __activemask(); stTriLanes++; if (lane == __ffs(am) - 1) stTriIters++; }
                        const float Oz = woopPlane(-origx, -origy, -origz, c00);
                        const float invDz = rcp_approx(woopDot(dirx, diry, dirz, c00));
                        const float t = fmulFtz(Oz, invDz);
                        const float Ox = woopPlane(origx, origy, origz, v11);
                        const float Dx = woopDot(dirx, diry, dirz, v11);
                        const float u = ffmaFtz(t, Dx, Ox);
                        const float Oy = woopPlane(origx, origy, origz, v22);
                        const float Dy = woopDot(dirx, diry, dirz, v22);
                        const float v = ffmaFtz(t, Dy, Oy);
                        if (woopAccept<WT>(t, tmin, hitT, u, v, v11, v22, WtRay{ origx, origy, origz, dirx, diry, dirz }, addr, hitIndex, p.triIndex)) {
                            hitT = t; hitIndex = addr;
                            if (WT) cullT = watertightCull(t);
                            if (BARY) { hitU = u; hitV = v; }
                            if (ANY_HIT) { nodeAddr = kSentinel; return true; }
                        }
                        return false;
                    };
                    if (UNROLL) {
                        // unrolled twice, the two row-0 register sets swapping roles: no copy of the requested row per triangle (4 of the
                        // loop's 41 instructions); C1 3 261 -> 3 319 Mrays/s, C2 within noise
                        float4 w00;
                        while (__float_as_int(v00.x) != (int)0x80000000) {
                            if (leafStep(triAddr, v00, w00)) break;
                            triAddr += 3;
                            if (__float_as_int(w00.x) == (int)0x80000000) break;
                            if (leafStep(triAddr, w00, v00)) break;
                            triAddr += 3;
                        }
                    } else
                    while (__float_as_int(v00.x) != (int)0x80000000) {
                        float4 n00;
                        if (leafStep(triAddr, v00, n00)) break;
                        v00 = n00; triAddr += 3;
                    }
                } else {
                    for (int triAddr = ~leafAddr;; triAddr += 3) {
                        const float4 v00 = ldg4(p.woop + triAddr), v11 = ldg4(p.woop + triAddr + 1), v22 = ldg4(p.woop + triAddr + 2);
                        if (__float_as_int(v00.x) == (int)0x80000000) break;       // leaf terminator
                        if (STATS) { const unsigned am = __activemask(); stTriLanes++; if (lane == __ffs(am) - 1) stTriIters++; }

                        const float Oz = woopPlane(-origx, -origy, -origz, v00);
                        const float invDz = rcp_approx(woopDot(dirx, diry, dirz, v00));
                        const float t = fmulFtz(Oz, invDz);
                        if (WT) {
                            const float Ox = woopPlane(origx, origy, origz, v11), Dx = woopDot(dirx, diry, dirz, v11);
                            const float Oy = woopPlane(origx, origy, origz, v22), Dy = woopDot(dirx, diry, dirz, v22);
                            const float u = ffmaFtz(t, Dx, Ox), v = ffmaFtz(t, Dy, Oy);
                            if (woopAccept<true>(t, tmin, hitT, u, v, v11, v22, WtRay{ origx, origy, origz, dirx, diry, dirz }, triAddr, hitIndex, p.triIndex)) {
                                hitT = t; hitIndex = triAddr; cullT = watertightCull(t);
                                if (BARY) { hitU = u; hitV = v; }
                                if (ANY_HIT) { nodeAddr = kSentinel; break; }
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
                                    hitT = t; hitIndex = triAddr;
                                    if (BARY) { hitU = u; hitV = v; }
                                    if (ANY_HIT) { nodeAddr = kSentinel; break; }
                                }
                            }
                        }
                    }
                }
                // a second leaf may be sitting in nodeAddr
                leafAddr = nodeAddr;
                if (nodeAddr < 0) nodeAddr = pop();
            }

            // dynamic fetch: too few lanes left -> leave and refill the idle ones
            if (__popc(__activemask()) < threshold) break;
            if (SHARE && !poolEmpty && *reinterpret_cast<volatile int*>(&s_share->idleWarps) > 0) break;
        }
    }

    if (SHARE) {
        // Every lane that worked on a shared ray of this CTA belongs to this CTA and has merged its hit.  Turn the merged
        // (Woop row, t) records into the final form: triangle id through triIndex, barycentrics recomputed with the
        // arithmetic of the leaf loop (same operands, same order: the same bits), t = tmax for a miss.
        __syncthreads();
        const int shared = s_share->listCount;
        for (int i = threadIdx.x; i < shared; i += BLOCK) {
            const int r = s_list[i];
            const unsigned long long key = __ldcg(reinterpret_cast<const unsigned long long*>(p.results + r));
            const int row = (int)(unsigned)(key & 0xffffffffull);
            const float4 d = ldg4(p.rays + 2 * (size_t)r + 1);
            int id = -1;
            float t = d.w, u = 0.f, v = 0.f;
            if (row != -1) {
                id = __ldg(p.triIndex + row);
                t = __uint_as_float((unsigned)(key >> 32));
                if (BARY) {
                    const float4 o = ldg4(p.rays + 2 * (size_t)r);
                    const float4 v11 = ldg4(p.woop + row + 1), v22 = ldg4(p.woop + row + 2);
                    u = ffmaFtz(t, woopDot(d.x, d.y, d.z, v11), woopPlane(o.x, o.y, o.z, v11));
                    v = ffmaFtz(t, woopDot(d.x, d.y, d.z, v22), woopPlane(o.x, o.y, o.z, v22));
                }
            }
            if (BARY) p.results[r] = make_int4(id, __float_as_int(t), __float_as_int(u), __float_as_int(v));
            else *reinterpret_cast<int2*>(p.results + r) = make_int2(id, __float_as_int(t));
        }
    }

    if (STATS) {
        const unsigned long long dbgIters = warpSum(stNodeIters) + (warpSum(stTriIters) << 32);
        // warp-reduce, one atomic per counter per warp
        for (int o = 16; o > 0; o >>= 1) {
            stNodeIters += __shfl_xor_sync(kFull, stNodeIters, o);
            stNodeLanes += __shfl_xor_sync(kFull, stNodeLanes, o);
            stTriIters += __shfl_xor_sync(kFull, stTriIters, o);
            stTriLanes += __shfl_xor_sync(kFull, stTriLanes, o);
            stFetches += __shfl_xor_sync(kFull, stFetches, o);
            stShares += __shfl_xor_sync(kFull, stShares, o);
        }
        if (lane == 0) {
            atomicAdd(p.stats + 0, stNodeIters); atomicAdd(p.stats + 1, stNodeLanes);
            atomicAdd(p.stats + 2, stTriIters);  atomicAdd(p.stats + 3, stTriLanes);
            atomicAdd(p.stats + 4, stFetches);
            // launch timeline (ns, global timer): first start, first warp to find the pool empty, mean and last exit
            const unsigned long long now = globalTimerNs();
            atomicMin(p.stats + 8, stStart);
            if (stPoolEmpty) atomicMin(p.stats + 9, stPoolEmpty);
            atomicAdd(p.stats + 10, now & 0xffffffffffull);     // low 40 bits: sums of <= 2^20 warps cannot overflow
            atomicMax(p.stats + 11, now);
            atomicAdd(p.stats + 12, 1ull);
            atomicAdd(p.stats + 13, stShares);
            const int gw = blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5);
            if (gw < kTimelineWarps) p.stats[16 + gw] = now;
            if (gw < kTimelineWarps) { p.stats[16 + kTimelineWarps + gw] = stLastFetch; p.stats[16 + 2 * kTimelineWarps + gw] = dbgIters - stItersAtFetch; }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// launcher
// ---------------------------------------------------------------------------------------------
struct Variant { int block; int stackSmem; int minBlocks; };

template <bool ANY_HIT, bool BARY, bool STATS, int STACK_SMEM, int BLOCK, int MIN_BLOCKS, int LEAF, bool WT = false>
static int launchOne(b2rt_ctx* ctx, TraceParams& p, cudaStream_t stream)
{
    auto kern = traceKernel<ANY_HIT, BARY, STATS, STACK_SMEM, BLOCK, MIN_BLOCKS, LEAF, WT>;
    // any-hit kernels carry the sharing queue behind the staged nodes; the small-CTA default gives up 40 nodes for it so
    // that four CTAs still fit the 164 KB shared-memory step (the next step, 196 KB, costs 32 KB of L1)
    constexpr size_t shareBytes = ANY_HIT ? sizeof(ShareArea) + (size_t)kShareQueue * sizeof(int4) : 0;
    int topNodes = ctx->cfg.top_nodes < 0 ? (BLOCK >= 1024 ? 2048 : (ANY_HIT ? 472 : 512)) : ctx->cfg.top_nodes;
    if (topNodes > ctx->topNodesAvail) topNodes = ctx->topNodesAvail;
    size_t stackBytes = (size_t)STACK_SMEM * BLOCK * sizeof(int) + shareBytes;
    // keep MIN_BLOCKS CTAs resident inside the 227 KB a CTA can opt into
    const size_t smemBudget = (size_t)(227 * 1024) / MIN_BLOCKS - 1024;
    topNodes &= ~7;                                  // planes stay 128-byte aligned
    while (topNodes > 0 && (size_t)topNodes * 56 + stackBytes > smemBudget) topNodes = (topNodes / 2) & ~7;
    p.topF4 = topNodes * 4;
    size_t smem = (size_t)topNodes * 56 + stackBytes;
    int blocksPerSm = 0;
    { const int rc = kernelOccupancy(ctx, kern, BLOCK, smem, &blocksPerSm); if (rc) return rc; }
    if (blocksPerSm < 1) return setError(ctx, B2RT_ERR_CUDA, "trace kernel does not fit on an SM");
    if (ctx->cfg.blocks_per_sm > 0 && ctx->cfg.blocks_per_sm < blocksPerSm) blocksPerSm = ctx->cfg.blocks_per_sm;
    long long grid = (long long)ctx->smCount * blocksPerSm;
    const long long needed = ((long long)p.numRays + BLOCK - 1) / BLOCK;
    if (grid > needed) grid = needed;
    if (grid < 1) grid = 1;
    const unsigned slot = ctx->launchSlot++ % b2rt_ctx::kLaunchSlots;
    p.warpCounter = ctx->warpCounter + slot * 32;
    if (ANY_HIT) {
        if (ctx->shareListCtas < (size_t)grid) {
            if (ctx->shareList) { cudaFree(ctx->shareList); ctx->shareList = nullptr; ctx->shareListCtas = 0; }
            B2RT_CUDA(ctx, cudaMalloc(&ctx->shareList, (size_t)b2rt_ctx::kLaunchSlots * grid * kShareList * sizeof(int)));
            ctx->shareListCtas = (size_t)grid;
        }
        p.shareList = ctx->shareList + (size_t)slot * ctx->shareListCtas * kShareList;
    }
    B2RT_CUDA(ctx, cudaMemsetAsync(p.warpCounter, 0, sizeof(int), stream));
    kern<<<(unsigned)grid, BLOCK, smem, stream>>>(p);
    B2RT_CHECK_LAUNCH(ctx, "traceKernel");
    return B2RT_OK;
}

template <int STACK_SMEM, int BLOCK, int MIN_BLOCKS, int LEAF = 0>
static int launchFlags(b2rt_ctx* ctx, TraceParams& p, unsigned flags, cudaStream_t stream)
{
    const bool any = (flags & B2RT_TRACE_ANY_HIT) != 0;
    const bool bary = (flags & B2RT_TRACE_WRITE_BARY) != 0;
    if (flags & B2RT_TRACE_WATERTIGHT)          // always writes barycentrics, never instrumented (two kernels per shape instead of six)
        return any ? launchOne<true, true, false, STACK_SMEM, BLOCK, MIN_BLOCKS, LEAF, true>(ctx, p, stream)
                   : launchOne<false, true, false, STACK_SMEM, BLOCK, MIN_BLOCKS, LEAF, true>(ctx, p, stream);
    if (flags & B2RT_TRACE_STATS) {
        // the instrumented build always writes barycentrics (one kernel fewer to carry)
        return any ? launchOne<true, true, true, STACK_SMEM, BLOCK, MIN_BLOCKS, LEAF>(ctx, p, stream)
                   : launchOne<false, true, true, STACK_SMEM, BLOCK, MIN_BLOCKS, LEAF>(ctx, p, stream);
    }
    if (any) return bary ? launchOne<true, true, false, STACK_SMEM, BLOCK, MIN_BLOCKS, LEAF>(ctx, p, stream)
                         : launchOne<true, false, false, STACK_SMEM, BLOCK, MIN_BLOCKS, LEAF>(ctx, p, stream);
    return bary ? launchOne<false, true, false, STACK_SMEM, BLOCK, MIN_BLOCKS, LEAF>(ctx, p, stream)
                : launchOne<false, false, false, STACK_SMEM, BLOCK, MIN_BLOCKS, LEAF>(ctx, p, stream);
}

int launchTrace(b2rt_ctx* ctx, const void* rays, void* results, int64_t numRays, unsigned flags,
                cudaStream_t stream)
{
    TraceParams p;
    p.rays = static_cast<const float4*>(rays);
    p.results = static_cast<int4*>(results);
    p.nodes = ctx->nodes;
    p.woop = ctx->woop;
    p.triIndex = ctx->triIndex;
    p.numRays = (int)numRays;
    p.topF4 = 0;
    p.fetchThreshold = ctx->cfg.fetch_threshold < 0 ? 20 : ctx->cfg.fetch_threshold;
    p.streamRays = (ctx->cfg.tune & B2RT_TUNE_NO_STREAM) ? 0 : 1;
    p.warpCounter = ctx->warpCounter;
    p.stats = ctx->statsDev;
    p.dummyLeaf = ~(int)(ctx->woopBytes / 16);       // first padding row: a terminator (api.cu:b2rt_set_bvh)
    p.shareList = nullptr;
    if (flags & B2RT_TRACE_STATS) {
        B2RT_CUDA(ctx, cudaMemsetAsync(ctx->statsDev, 0, 16 * sizeof(unsigned long long), stream));
        B2RT_CUDA(ctx, cudaMemsetAsync(ctx->statsDev + 8, 0xff, 2 * sizeof(unsigned long long), stream));    // minima
    }
    int variant = ctx->cfg.variant;
    if (variant < 0) {
        // Measured on B200 (profiles/r01_explore_*_final.log, r01_explore_c{2,3,5}_wide_policy.log).  Binary kernel: one
        // 1024-thread CTA per SM with a 2048-node shared top of the tree and the look-ahead leaf loop wins on frame-sized
        // batches (shorter tail, more node visits served from shared memory); any-hit batches share subtrees inside a CTA
        // once the pool is dry, two 512-thread CTAs per SM; four 256-thread CTAs per SM on multi-million-ray batches.
        variant = (flags & B2RT_TRACE_ANY_HIT) ? 9 : 8;
        if (numRays > (4 << 20)) variant = 1;
        // The 8-wide compressed private form (trace_wide.cu; built on first use, api.cu:ensureWide) wins once the BVH is
        // large against what an SM keeps close -- 80 bytes per 8 children instead of 64 per 2, fewer dependent fetches per
        // ray: knot 910 K triangles (70 MB of nodes + Woop rows): AO any-hit 3 100 vs 2 900, 64 Mi diffuse rays 3 210 vs
        // 3 020; strands 6.5 M triangles (1.2 GB): 574 vs 425.  On the 351 K-triangle interior (32 MB) the binary kernel
        // keeps a 25 % lead (4 800 vs 3 780), so the switch is the size of the two arrays.
        if (ctx->nodeBytes + ctx->woopBytes > (size_t(48) << 20)) {
            // ... unless the collapsed tree is too deep for that kernel's stack (a degenerate caller-built BVH): stay binary
            const int rc = ensureWide(ctx);
            if (rc) return rc;
            // block claims (variant 114) above 12 Mi rays: a warp can hold 127 unstarted rays when the pool runs dry, 7 % of an
            // 8 Mi-ray batch over 4 736 warps -- measured on shards of the C5 batch with the kernels as shipped (108 vs 114): 8 Mi rays
            // 3 419 vs 3 310, 16 Mi 3 520 vs 3 570, 21 Mi 3 536 vs 3 605, 32 Mi 3 598 vs 3 725 Mrays/s (the whole 64 Mi batch: 3 802 with 114)
            // (profiles/r02_explore_c5_last_session.log; before block classification the crossover was at 24 Mi:
            // r02_explore_c5_shards.log)
            if (ctx->wideMaxDepth <= kWideMaxDepth) variant = (numRays > (12 << 20)) ? 114 : 108;
        }
    }
    if (variant >= 100) return launchTraceWide(ctx, rays, results, numRays, flags, variant, stream);
    switch (variant) {
    default:
    case 1: return launchFlags<12, 256, 4>(ctx, p, flags, stream);      // four 256-thread CTAs per SM (<= 64 regs), reference-shaped leaf loop
    case 8: return launchFlags<12, 1024, 1, 2>(ctx, p, flags, stream);  // one 1024-thread CTA per SM, 2048-node shared top of the tree, look-ahead leaf loop unrolled twice
    case 9: return launchFlags<12, 512, 2, 2>(ctx, p, flags, stream);   // two 512-thread CTAs per SM, the same leaf loop (any-hit batches)
    case 10: return launchFlags<12, 256, 4, 2>(ctx, p, flags, stream);  // variant 1 with that leaf loop
    case 11: return launchFlags<12, 256, 4, 1>(ctx, p, flags, stream);  // 10 with the look-ahead leaf loop not unrolled (LEAF == 1): the bit-identity test's third shape
#ifdef B2RT_EXPLORE
    case 18: return launchFlags<12, 1024, 1, 1>(ctx, p, flags, stream); // 8 not unrolled: C1 3 319 -> 3 261, C2 within noise (profiles/r02_explore_c5_last_session.log)
    case 19: return launchFlags<12, 512, 2, 1>(ctx, p, flags, stream);  // 9 not unrolled
#endif
#ifdef B2RT_EXPLORE
    // shapes measured in round 1 and not shipped; build with -DB2RT_EXPLORE to get them back
    case 0: return launchFlags<12, 256, 5>(ctx, p, flags, stream);   // <= 48 regs, 40 warps/SM
    case 2: return launchFlags<0, 256, 5>(ctx, p, flags, stream);    // stack entirely in local memory
    case 3: return launchFlags<16, 128, 8>(ctx, p, flags, stream);   // small CTAs
    case 6: return launchFlags<12, 1024, 1>(ctx, p, flags, stream);  // variant 8 with the reference-shaped leaf loop
    case 7: return launchFlags<12, 512, 2>(ctx, p, flags, stream);   // variant 9 with the reference-shaped leaf loop
#endif
    }
}


// ---------------------------------------------------------------------------------------------
// canonical work counters (b2rt_count_work): ordered stack traversal without speculation or
// postponement (kepler_dynamic_fetch.cu:204-408 reduced to its scalar algorithm), one thread per
// ray, IEEE division -- counts, does not store hits
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) countWorkKernel(const float4* __restrict__ rays, int numRays, int anyHit,
                                                       const float4* __restrict__ nodes, const float4* __restrict__ woop,
                                                       unsigned long long* out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long cN = 0, cT = 0, cL = 0, cU = 0, cV = 0, cH = 0, cS = 0;
    if (i < numRays) {
        const float4 o = __ldg(rays + 2 * (size_t)i), d = __ldg(rays + 2 * (size_t)i + 1);
        const float tmin = o.w;
        float hitT = d.w;
        const float ooeps = 8.271806125530277e-25f;
        const float idirx = __fdiv_rn(1.0f, fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x));
        const float idiry = __fdiv_rn(1.0f, fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y));
        const float idirz = __fdiv_rn(1.0f, fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z));
        const float oodx = __fmul_rn(o.x, idirx), oody = __fmul_rn(o.y, idiry), oodz = __fmul_rn(o.z, idirz);
        int stack[72];
        int sp = 0, nodeAddr = 0, hitIndex = -1;
        stack[0] = kSentinel;
        while (nodeAddr != kSentinel) {
            if (nodeAddr >= 0) {
                const float4 n0xy = ldg4(nodes + nodeAddr), n1xy = ldg4(nodes + nodeAddr + 1), nz = ldg4(nodes + nodeAddr + 2), cn = ldg4(nodes + nodeAddr + 3);
                cN++;
                const float c0lox = __fmaf_rn(n0xy.x, idirx, -oodx), c0hix = __fmaf_rn(n0xy.y, idirx, -oodx);
                const float c0loy = __fmaf_rn(n0xy.z, idiry, -oody), c0hiy = __fmaf_rn(n0xy.w, idiry, -oody);
                const float c0loz = __fmaf_rn(nz.x, idirz, -oodz), c0hiz = __fmaf_rn(nz.y, idirz, -oodz);
                const float c1loz = __fmaf_rn(nz.z, idirz, -oodz), c1hiz = __fmaf_rn(nz.w, idirz, -oodz);
                const float c1lox = __fmaf_rn(n1xy.x, idirx, -oodx), c1hix = __fmaf_rn(n1xy.y, idirx, -oodx);
                const float c1loy = __fmaf_rn(n1xy.z, idiry, -oody), c1hiy = __fmaf_rn(n1xy.w, idiry, -oody);
                const float c0min = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), tmin));
                const float c0max = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), hitT));
                const float c1min = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), tmin));
                const float c1max = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), hitT));
                const bool t0 = c0max >= c0min, t1 = c1max >= c1min;
                const int c0 = __float_as_int(cn.x), c1 = __float_as_int(cn.y);
                if (!t0 && !t1) nodeAddr = stack[sp--];
                else {
                    nodeAddr = t0 ? c0 : c1;
                    if (t0 && t1) {
                        int farChild = c1;
                        if (c1min < c0min) { farChild = nodeAddr; nodeAddr = c1; }
                        stack[++sp] = farChild;
                        if ((unsigned long long)sp > cS) cS = sp;
                    }
                }
            } else {
                cL++;
                bool done = false;
                for (int triAddr = ~nodeAddr;; triAddr += 3) {
                    const float4 v00 = ldg4(woop + triAddr);
                    if (__float_as_int(v00.x) == (int)0x80000000) break;
                    const float4 v11 = ldg4(woop + triAddr + 1), v22 = ldg4(woop + triAddr + 2);
                    cT++;
                    const float Oz = __fmaf_rn(-o.z, v00.z, __fmaf_rn(-o.y, v00.y, __fmaf_rn(-o.x, v00.x, v00.w)));
                    const float invDz = __fdiv_rn(1.0f, __fmaf_rn(d.z, v00.z, __fmaf_rn(d.y, v00.y, __fmul_rn(d.x, v00.x))));
                    const float t = fmulFtz(Oz, invDz);
                    if (t > tmin && t < hitT) {
                        cU++;
                        const float Ox = __fmaf_rn(o.z, v11.z, __fmaf_rn(o.y, v11.y, __fmaf_rn(o.x, v11.x, v11.w)));
                        const float Dx = __fmaf_rn(d.z, v11.z, __fmaf_rn(d.y, v11.y, __fmul_rn(d.x, v11.x)));
                        const float u = ffmaFtz(t, Dx, Ox);
                        if (u >= 0.0f) {
                            cV++;
                            const float Oy = __fmaf_rn(o.z, v22.z, __fmaf_rn(o.y, v22.y, __fmaf_rn(o.x, v22.x, v22.w)));
                            const float Dy = __fmaf_rn(d.z, v22.z, __fmaf_rn(d.y, v22.y, __fmul_rn(d.x, v22.x)));
                            const float v = ffmaFtz(t, Dy, Oy);
                            if (v >= 0.0f && faddFtz(u, v) <= 1.0f) {
                                hitT = t; hitIndex = triAddr;
                                if (anyHit) { done = true; break; }
                            }
                        }
                    }
                }
                nodeAddr = done ? kSentinel : stack[sp--];
            }
        }
        cH = (hitIndex != -1) ? 1 : 0;
    }
    unsigned long long v[7] = { cN, cT, cL, cU, cV, cH, 0 };
    for (int k = 0; k < 6; k++) {
        for (int s = 16; s > 0; s >>= 1) v[k] += __shfl_xor_sync(kFull, v[k], s);
        if ((threadIdx.x & 31) == 0 && v[k]) atomicAdd(out + 1 + k, v[k]);
    }
    for (int s = 16; s > 0; s >>= 1) { const unsigned long long t = __shfl_xor_sync(kFull, cS, s); cS = t > cS ? t : cS; }
    if ((threadIdx.x & 31) == 0) atomicMax(out + 7, cS);
    unsigned long long cM = cN + cT;                // the longest dependent chain of one ray: node visits + triangle tests
    for (int s = 16; s > 0; s >>= 1) { const unsigned long long t = __shfl_xor_sync(kFull, cM, s); cM = t > cM ? t : cM; }
    if ((threadIdx.x & 31) == 0) atomicMax(out + 0, cM);
}

int countWork(b2rt_ctx* ctx, const void* rays, int64_t numRays, int anyHit, b2rt_work_counters* out, cudaStream_t stream)
{
    B2RT_CUDA(ctx, cudaMemsetAsync(ctx->statsDev, 0, 16 * sizeof(unsigned long long), stream));
    countWorkKernel<<<(unsigned)((numRays + 255) / 256), 256, 0, stream>>>(static_cast<const float4*>(rays), (int)numRays, anyHit,
                                                                          ctx->nodes, ctx->woop, ctx->statsDev);
    B2RT_CHECK_LAUNCH(ctx, "countWorkKernel");
    unsigned long long h[8];
    B2RT_CUDA(ctx, cudaMemcpyAsync(h, ctx->statsDev, sizeof(h), cudaMemcpyDeviceToHost, stream));
    B2RT_CUDA(ctx, cudaStreamSynchronize(stream));
    out->rays = (uint64_t)numRays; out->max_chain = h[0]; out->nodes = h[1]; out->tris = h[2]; out->leaves = h[3];
    out->ustage = h[4]; out->vstage = h[5]; out->hits = h[6]; out->max_stack = h[7];
    return B2RT_OK;
}

}  // namespace b2rt
