/*
 * stubs.c -- link-time stand-ins so the reference's host code runs on a machine with no GPU
 * and no graphics stack.  TEST INFRASTRUCTURE ONLY (oracle/_ref/ref_cpu).
 *
 * CUDA runtime: "device" memory is host memory (cudaMalloc = malloc, cudaMemcpy = memcpy), so the
 * reference's RayBuffer / PixelTable / CudaBVH / RayGen host logic runs unchanged and the CPU
 * launchers in ref_driver.cc can read and write the same pointers.
 * CUDA driver API and GL buffer objects: named by the reference's Buffer.cc but never reached
 * (only Buffer::CPU ownership is exercised); they abort if they ever are.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stddef.h>

typedef int cudaError_t;
enum { b2o_cudaSuccess = 0 };

#ifndef B2O_REAL_CUDA      /* oracle/_ref/route_a links the real CUDA runtime and driver: only the GL names below stay stubs */
cudaError_t cudaMalloc(void** p, size_t n) { *p = malloc(n ? n : 1); return *p ? 0 : 2; }
cudaError_t cudaFree(void* p) { free(p); return 0; }
cudaError_t cudaMemcpy(void* d, const void* s, size_t n, int kind) { (void)kind; memmove(d, s, n); return 0; }
cudaError_t cudaDeviceSynchronize(void) { return 0; }
const char* cudaGetErrorString(cudaError_t e) { (void)e; return "b2rt oracle stub: CUDA runtime is not linked"; }
#endif

#define UNREACHED(name) \
    int name() { fprintf(stderr, "b2rt oracle stub: " #name " reached\n"); abort(); return 0; }

UNREACHED(cuGLMapBufferObject_v2)
UNREACHED(cuGLRegisterBufferObject)
UNREACHED(cuGLUnmapBufferObject)
UNREACHED(cuGLUnregisterBufferObject)
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemAllocHost_v2)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemAlloc_v2)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemFreeHost)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemFree_v2)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemcpyDtoDAsync_v2)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemcpyDtoD_v2)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemcpyDtoHAsync_v2)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemcpyDtoH_v2)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemcpyHtoDAsync_v2)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemcpyHtoD_v2)
#endif
#ifndef B2O_REAL_CUDA
UNREACHED(cuMemsetD8_v2)
#endif
UNREACHED(glBindBuffer)
UNREACHED(glBufferData)
UNREACHED(glBufferSubData)
UNREACHED(glDeleteBuffers)
UNREACHED(glGenBuffers)
UNREACHED(glGetBufferParameteriv)
UNREACHED(glGetBufferSubData)
UNREACHED(glGetIntegerv)
#ifndef B2O_REAL_CUDA
UNREACHED(cuInit)
#endif
