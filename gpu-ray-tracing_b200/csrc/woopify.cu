// woopify.cu -- the Woop rows of a Compact2 BVH computed on the device.
//
// Replaces the per-triangle part of CudaBVH::createCompact (rt/cuda/CudaBVH.cc:270-358): woopifyTri (CudaBVH.cc:361-380)
// builds the matrix [v0-v2 | v1-v2 | (v0-v2)x(v1-v2) | v2], inverts it with the cofactor expansion of Math.hh:935-983 and
// stores rows 2, 0, 1 of the inverse.  For 6.5 M triangles with SBVH duplication that is 20 M inversions on one host
// core in the reference; here one thread per triangle REFERENCE does it.  The result has to be bit-identical to the host
// computation (the rows are the input of the intersection test, so any difference moves hit distances): every product,
// sum and the one division are spelled out with the round-to-nearest intrinsics in the order the host code evaluates
// them, so nothing is contracted into FMAs or re-associated.
//
// Layout contract (CudaBVH.hh:47-50): three 16-byte rows per triangle reference, one terminator row (x = 0x80000000)
// closing each leaf.  The caller passes, for each reference, its first row and its triangle; every row that no reference
// covers is written as a terminator.
#include "common.cuh"

namespace b2rt {

__device__ __forceinline__ float mulr(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float addr(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float subr(float a, float b) { return __fsub_rn(a, b); }

// 3x3 determinant, operation order of Math.hh:935-940 (left to right)
__device__ __forceinline__ float det3(float s00, float s01, float s02, float s10, float s11, float s12, float s20, float s21, float s22)
{
    float r = mulr(mulr(s00, s11), s22);
    r = subr(r, mulr(mulr(s00, s12), s21));
    r = addr(r, mulr(mulr(s10, s21), s02));
    r = subr(r, mulr(mulr(s10, s22), s01));
    r = addr(r, mulr(mulr(s20, s01), s12));
    r = subr(r, mulr(mulr(s20, s02), s11));
    return r;
}

__global__ void __launch_bounds__(256) fillTerminatorsKernel(uint4* __restrict__ woop, long long numRows)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < numRows) woop[i] = make_uint4(0x80000000u, 0x80000000u, 0x80000000u, 0x80000000u);
}

__global__ void __launch_bounds__(128) woopifyKernel(const float* __restrict__ verts, const int* __restrict__ vtxIndex,
                                                     const int* __restrict__ refRow, const int* __restrict__ refTri,
                                                     long long numRefs, float4* __restrict__ woop)
{
    const long long ref = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (ref >= numRefs) return;
    const int tri = __ldg(refTri + ref);
    const int i0 = __ldg(vtxIndex + 3 * (size_t)tri), i1 = __ldg(vtxIndex + 3 * (size_t)tri + 1), i2 = __ldg(vtxIndex + 3 * (size_t)tri + 2);
    float v0[3], v1[3], v2[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        v0[k] = __ldg(verts + 3 * (size_t)i0 + k); v1[k] = __ldg(verts + 3 * (size_t)i1 + k); v2[k] = __ldg(verts + 3 * (size_t)i2 + k);
    }
    float m[4][4];      // m[row][column]
    float e0[3], e1[3];
#pragma unroll
    for (int k = 0; k < 3; k++) { e0[k] = subr(v0[k], v2[k]); e1[k] = subr(v1[k], v2[k]); }
    const float nn[3] = { subr(mulr(e0[1], e1[2]), mulr(e0[2], e1[1])), subr(mulr(e0[2], e1[0]), mulr(e0[0], e1[2])),
                          subr(mulr(e0[0], e1[1]), mulr(e0[1], e1[0])) };
#pragma unroll
    for (int r = 0; r < 3; r++) { m[r][0] = e0[r]; m[r][1] = e1[r]; m[r][2] = nn[r]; m[r][3] = v2[r]; }
    m[3][0] = m[3][1] = m[3][2] = 0.0f; m[3][3] = 1.0f;

    float inv[4][4];
    float d = 0.0f;
#pragma unroll
    for (int i = 0; i < 4; i++) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            // minor: rows != j, columns != i
            float s[3][3];
#pragma unroll
            for (int k = 0; k < 3; k++)
#pragma unroll
                for (int l = 0; l < 3; l++) s[k][l] = m[(k < j) ? k : k + 1][(l < i) ? l : l + 1];
            const float sign = ((i + j) & 1) ? -1.0f : 1.0f;
            const float dd = mulr(det3(s[0][0], s[0][1], s[0][2], s[1][0], s[1][1], s[1][2], s[2][0], s[2][1], s[2][2]), sign);
            inv[i][j] = dd;
            d = addr(d, mulr(dd, m[j][i]));
        }
    }
    const float rd = (d != 0.0f) ? __fdiv_rn(1.0f, d) : 0.0f;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) inv[i][j] = mulr(mulr(inv[i][j], rd), 4.0f);

    float x = inv[2][0];
    if (x == 0.0f) x = 0.0f;            // -0.0f would read as the leaf terminator (CudaBVH.cc:316-317)
    const int row = __ldg(refRow + ref);
    woop[row] = make_float4(x, inv[2][1], inv[2][2], -inv[2][3]);
    woop[row + 1] = make_float4(inv[0][0], inv[0][1], inv[0][2], inv[0][3]);
    woop[row + 2] = make_float4(inv[1][0], inv[1][1], inv[1][2], inv[1][3]);
}

}  // namespace b2rt

using namespace b2rt;

extern "C" int b2rt_woopify(b2rt_ctx* ctx, const float* verts, const int32_t* vtxIndex, const int32_t* refRow, const int32_t* refTri,
                            int64_t numRefs, void* woop, int64_t numRows, void* stream)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (numRefs < 0 || numRows < 0 || numRows > 0x7fffffff || (numRows > 0 && !woop) ||
        (numRefs > 0 && (!verts || !vtxIndex || !refRow || !refTri)))
        return setError(ctx, B2RT_ERR_ARG, "b2rt_woopify: bad arguments (refs=%lld rows=%lld)", (long long)numRefs, (long long)numRows);
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (numRows > 0) {
        fillTerminatorsKernel<<<(unsigned)((numRows + 255) / 256), 256, 0, s>>>(static_cast<uint4*>(woop), numRows);
        B2RT_CHECK_LAUNCH(ctx, "fillTerminatorsKernel");
    }
    if (numRefs > 0) {
        woopifyKernel<<<(unsigned)((numRefs + 127) / 128), 128, 0, s>>>(verts, vtxIndex, refRow, refTri, numRefs, static_cast<float4*>(woop));
        B2RT_CHECK_LAUNCH(ctx, "woopifyKernel");
    }
    return B2RT_OK;
}
