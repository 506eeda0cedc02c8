/*
 * host_emul.h -- force-included (g++ -include) in front of the reference's auxiliary CUDA kernels
 * (src/rt/ray/RayGenKernels.cu, src/rt/ray/RayBufferKernels.cu, src/rt/cuda/RendererKernels.cu) so
 * that their kernel BODIES compile as plain host C++ and run serially, one emulated thread at a
 * time, inside oracle/_ref/ref_cpu.  TEST INFRASTRUCTURE ONLY; not product code.
 *
 * That makes the reference's own code -- not our restatement -- the generator of primary / AO /
 * diffuse rays, Morton keys, the reorder gather and the reconstructed pixels on a machine with no
 * GPU.  Only kernels whose threads do not cooperate are run this way; findAABBKernel and
 * countHitsKernel reduce through shared memory with __syncthreads and warp-synchronous steps
 * (RayBufferKernels.cu:106-125, RendererKernels.cu:136-154), which a serial loop cannot emulate, so
 * their launchers are renamed away by oracle/Makefile and stay bound to b2rt_oracle.c.
 * Two sed edits are applied to the copies (never to the repository): `k<<<grid,block>>>()` becomes
 * b2o_emul_launch(k, grid, block), and genMortonKeysKernel's hashbuf[32*4*6] becomes [32*8*6]
 * (CudaKernel::setGrid forces 32x8 blocks, CudaKernel.cc:64, so the shipped size overruns).
 * FW_CUDA is 0 while Defs.hh / Math.hh / Util.hh are parsed (host flavour of the math library) and
 * 1 afterwards, so the *.hh kernel headers declare their __constant__ parameter blocks.
 * __sinf/__cosf resolve to the host's sinf/cosf (the GPU build uses the fast-math intrinsics).
 */
#ifndef B2RT_ORACLE_HOST_EMUL_H
#define B2RT_ORACLE_HOST_EMUL_H
#include <cuda_runtime.h>
#include "base/Defs.hh"
#include "base/Math.hh"
#include "Util.hh"
#undef FW_CUDA
#define FW_CUDA 1
#define globalThreadIdx (threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * (blockIdx.x + gridDim.x * blockIdx.y)))
#include <math.h>
#include <string.h>
#undef __global__
#undef __device__
#undef __constant__
#undef __shared__
#undef __inline__
#define __global__
#define __device__
#define __constant__
#define __shared__ static
#define __inline__ inline
static uint3 threadIdx, blockIdx;
static dim3 blockDim, gridDim;
static inline void __syncthreads(void) {}
static inline float __int_as_float(int v) { float f; memcpy(&f, &v, 4); return f; }
static inline int __float_as_int(float f) { int v; memcpy(&v, &f, 4); return v; }
static inline unsigned atomicAdd(unsigned* p, unsigned v) { unsigned o = *p; *p = o + v; return o; }
static inline int atomicAdd(int* p, int v) { int o = *p; *p = o + v; return o; }
static inline unsigned atomicCAS(unsigned* p, unsigned c, unsigned v) { unsigned o = *p; if (o == c) *p = v; return o; }
template <class K> static inline void b2o_emul_launch(K kernel, dim3 grid, dim3 block)
{
    gridDim = grid; blockDim = block;
    for (unsigned by = 0; by < grid.y; by++) for (unsigned bx = 0; bx < grid.x; bx++)
    for (unsigned ty = 0; ty < block.y; ty++) for (unsigned tx = 0; tx < block.x; tx++) {
        blockIdx.x = bx; blockIdx.y = by; blockIdx.z = 0; threadIdx.x = tx; threadIdx.y = ty; threadIdx.z = 0;
        kernel();
    }
}
template <class T> static inline cudaError_t cudaMemcpyToSymbol(T& sym, const void* src, size_t n, size_t off = 0, cudaMemcpyKind = cudaMemcpyHostToDevice)
{ memcpy((char*)&sym + off, src, n); return cudaSuccess; }
template <class T> static inline cudaError_t cudaMemcpyFromSymbol(void* dst, T& sym, size_t n, size_t off = 0, cudaMemcpyKind = cudaMemcpyDeviceToHost)
{ memcpy(dst, (const char*)&sym + off, n); return cudaSuccess; }

#endif
