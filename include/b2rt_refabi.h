/*
 * b2rt_refabi.h -- the reference's own kernel-launcher ABI, served by the Blackwell-native core.
 *
 * The reference links exactly one src/rt/kernels/*.cu plus RayGenKernels.cu, RayBufferKernels.cu
 * and RendererKernels.cu (Makefile:73-76); its host classes (CudaTracer.cc, RayGen.cc, RayBuffer.cc,
 * Renderer.cc) reach the GPU only through the C-linkage launchers those files define.
 * libb2rt_refabi.so defines the SAME symbols with the same argument layouts on top of libb2rt.so
 * (include/b2rt.h), so a reference build that drops those four .cu files from its link line and adds
 * `-lb2rt_refabi -lb2rt` runs CudaTracer::traceBatch(RayBuffer&), RayGen::primary/ao,
 * RayBuffer::mortonSort and Renderer::getTotalNumRays/updateResult unchanged on the new kernels.
 * See INTEGRATION.md for the Makefile diff.
 *
 * The reference declares these `extern "C"` but passes C++ references (Vec2i&, XxxInput&); at the
 * ABI a reference is a pointer, so they are spelled as pointers here.  The parameter blocks below
 * restate the reference's POD layouts field for field (x86-64 SysV); FW::Vec3f = 3 floats,
 * FW::Mat4f = 16 floats column-major, FW::Vec2i = 2 ints.
 *
 * Error convention: these entry points keep the reference's (cutil_inline_runtime.h:33-42,
 * Defs.cc:41-54): print the message and exit(-1).  The int-returning b2rt_* layer underneath never
 * exits.  State: one process-global b2rt_ctx per CUDA device, created on first use (the reference
 * keeps process-global texture bindings and __constant__ blocks in the same way).
 */
#ifndef B2RT_REFABI_H
#define B2RT_REFABI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ray/RayGenKernels.hh:40-51 */
typedef struct {
    float    origin[3];
    float    nscreenToWorld[16];
    int32_t  w, h;
    float    maxDist;
    void*    rays;           /* Ray*        (device) */
    int32_t* idToSlot;
    int32_t* slotToID;
    int32_t* indexToPixel;   /* const S32*  (device) */
} b2rt_ref_RayGenPrimaryInput;

/* ray/RayGenKernels.hh:55-68 */
typedef struct {
    int32_t  firstInputSlot, numInputRays, numSamples;
    float    maxDist;
    uint32_t randomSeed;
    void*    inRays;         /* const Ray* */
    void*    inResults;      /* const RayResult* */
    void*    outRays;
    int32_t* outIDToSlot;
    int32_t* outSlotToID;
    void*    normals;        /* const Vec3f*, one per triangle */
} b2rt_ref_RayGenAOInput;

/* ray/RayBufferKernels.hh:54-91 */
typedef struct { int32_t numRays, raysPerThread; void* inRays; } b2rt_ref_FindAABBInput;
typedef struct { float aabbLo[3], aabbHi[3]; } b2rt_ref_FindAABBOutput;
typedef struct { int32_t numRays; float aabbLo[3], aabbHi[3]; void* inRays; void* outKeys; } b2rt_ref_GenMortonKeysInput;
typedef struct {
    int32_t  numRays;
    void*    inKeys;         /* const MortonKey* (28 B) */
    void*    inRays;
    int32_t* inSlotToID;
    void*    outRays;
    int32_t* outIDToSlot;
    int32_t* outSlotToID;
} b2rt_ref_ReorderRaysInput;

/* cuda/RendererKernels.hh:54-82 */
typedef struct {
    int32_t   numRaysPerPrimary, firstPrimary, numPrimary;
    uint8_t   isPrimary, isAO, isDiffuse;     /* C++ bool */
    int32_t*  primarySlotToID;
    void*     primaryResults;
    int32_t*  batchIDToSlot;
    void*     batchResults;
    uint32_t* triMaterialColor;
    uint32_t* triShadedColor;
    uint32_t* pixels;
} b2rt_ref_ReconstructInput;
typedef struct { int32_t numRays; void* rayResults; int32_t raysPerThread; } b2rt_ref_CountHitsInput;

/* kernels/CudaTracerKernels.hh:42-52, defined by kernels/kepler_dynamic_fetch.cu:417-474 */
void  bind_CudaBVHTexture(void* nodes /* float4* */, int64_t nodeBytes, void* woop /* float4* */, int64_t woopBytes,
                          int* triIndices, int64_t triIndexBytes);
void  unbind_CudaBVHTexture(void);
/* returns the kernel time in MILLISECONDS (cudaEventElapsedTime, kepler_dynamic_fetch.cu:471-473).
 * threads / blockSize are the reference's launch-shape request (CudaTracer.cc:152-161); the new
 * kernel sizes its own persistent grid from the SM count and ignores them.  Only nodesA, trisA and
 * triIndices are read (BVHLayout_Compact2, CudaBVH.cc:120-140). */
float launch_tracingKernel(int32_t threads, int* blockSize /* Vec2i& */, int numRays, bool anyHit, void* rays /* float4* */,
                           void* results /* int4* */, void* nodesA, void* nodesB, void* nodesC, void* nodesD, void* trisA,
                           void* trisB, void* trisC, int* triIndices);
void  copy_tracing_results(void* hostResults /* RayResult* */, void* devResults, int32_t n);

/* ray/RayGenKernels.hh:107-111 */
void launch_rayGenPrimaryKernel(int32_t nthreads, b2rt_ref_RayGenPrimaryInput* in);
void launch_rayGenAOKernel(int32_t nthreads, b2rt_ref_RayGenAOInput* in);

/* ray/RayBufferKernels.hh:112-117 */
void launch_findAABBKernel(int32_t nthreads, const int* blockSize, b2rt_ref_FindAABBInput* in, b2rt_ref_FindAABBOutput* out);
void launch_genMortonKeysKernel(int32_t nthreads, const int* blockSize, b2rt_ref_GenMortonKeysInput* in);
void launch_reorderRaysKernel(int32_t nthreads, b2rt_ref_ReorderRaysInput* in);

/* cuda/RendererKernels.hh:88-92 */
void    launch_reconstructKernel(int32_t nthreads, b2rt_ref_ReconstructInput* in);
int32_t launch_countHitsKernel(int32_t nthreads, const int* blockSize, b2rt_ref_CountHitsInput* in);

#ifdef __cplusplus
}
#endif
#endif /* B2RT_REFABI_H */
