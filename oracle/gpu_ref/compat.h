/*
 * compat.h -- force-included (nvcc -include) in front of the reference's CUDA tracer so that
 * src/rt/kernels/kepler_dynamic_fetch.cu builds with CUDA 12.9 for sm_100a.
 * TEST INFRASTRUCTURE / TIMED BASELINE ONLY (oracle/_ref/libref_gpu.so); not product code.
 *
 * The kernel is written against APIs CUDA 12 removed (SURVEY.md probe table):
 *   texture<T,1> references + tex1Dfetch(texref, i) + cudaBindTexture   -> a __device__ pointer
 *       wrapper read with __ldg (same read-only data path on sm_100)
 *   __ballot(p)                                                          -> __ballot_sync(__activemask(), p)
 * Two more edits cannot be made from a header and are applied by oracle/Makefile with sed while
 * copying the .cu into oracle/_ref/ (never into the repository):
 *   vote.ballot.b32 (inline PTX, :304-310)  -> activemask + vote.sync.ballot.b32
 *   nextRayArray[MaxBlockHeight] (:96)      -> nextRayArray[8]  (blockDim.y is forced to 8 by
 *       CudaKernel::setGrid, CudaKernel.cc:64, so the shipped 6 overruns shared memory)
 * Everything else -- traversal order, postponement, dynamic fetch threshold, video-instruction
 * min/max, local-memory stack, fast-math flags -- is the reference's code, unmodified.
 */
#ifndef B2RT_ORACLE_GPU_COMPAT_H
#define B2RT_ORACLE_GPU_COMPAT_H

#include <cuda_runtime.h>

template <class T, int D>
struct b2o_texref { const T* ptr; };

template <class T>
__device__ __forceinline__ T tex1Dfetch(const b2o_texref<T, 1>& t, int i) { return __ldg(t.ptr + i); }

template <class T>
inline cudaError_t cudaBindTexture(size_t* offset, b2o_texref<T, 1>& tex, const void* ptr, cudaChannelFormatDesc, size_t)
{
    if (offset) *offset = 0;
    const T* p = static_cast<const T*>(ptr);
    return cudaMemcpyToSymbol(tex, &p, sizeof(p));
}
template <class T>
inline cudaError_t cudaUnbindTexture(b2o_texref<T, 1>&) { return cudaSuccess; }

#ifdef B2O_STATIC_TEXREF   /* several translation units include the declarations (route A: the reference's fetch_*.cu) */
#define texture static __device__ ::b2o_texref
#else
#define texture __device__ ::b2o_texref
#endif
#define __ballot(p) __ballot_sync(__activemask(), (p))

#endif
