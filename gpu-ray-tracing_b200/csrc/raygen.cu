// raygen.cu -- primary and AO / diffuse ray generation.
//
// Replaces ray/RayGenKernels.cu:79-113 (rayGenPrimaryKernel), :117-227 (rayGenAOKernel) and the
// launchers :295-330.  Compiled with -fmad=false and IEEE division / sqrt so that every value
// that does not go through sin/cos is bit-identical to the CPU oracle (and to the reference's
// host build); the reference's device build uses --use_fast_math (Makefile:120).
//
// B200 shape: both kernels are pure streaming writes (40 B per ray: 32 B ray + two 4 B maps).
//  * primary: one thread per slot, two 16 B stores per ray.
//  * AO/diffuse: ONE THREAD PER OUTPUT RAY instead of the reference's one thread per primary
//    with a strided inner loop over samples (RayGenKernels.cu:178-226, whose stores are
//    numSamples*32 B apart across a warp): consecutive lanes write consecutive rays, so every
//    store instruction of a warp covers contiguous 512 B.  The per-primary frame is recomputed
//    per sample; that arithmetic is free next to the HBM writes.
//  * parameters travel as kernel arguments, not through cudaMemcpyToSymbol + a device sync.
#include "common.cuh"

namespace b2rt {

struct PrimaryParams {
    float origin[3];
    float m[16];            // nscreenToWorld, column-major (FW::Mat4f)
    int w, h;
    float maxDist;
    const int32_t* indexToPixel;
    float4* rays;
    int32_t* idToSlot;
    int32_t* slotToID;
};

__device__ __forceinline__ float rcpSafe(float a) { return (a != 0.0f) ? 1.0f / a : 0.0f; }   // Math.hh:113

__device__ __forceinline__ void normalize3(float& x, float& y, float& z)
{
    float l = 0.0f;
    l += x * x; l += y * y; l += z * z;
    const float s = 1.0f * rcpSafe(sqrtf(l));       // Math.hh:142
    x *= s; y *= s; z *= s;
}

__global__ void __launch_bounds__(256) rayGenPrimaryKernel(const PrimaryParams p)
{
    const int taskIdx = blockIdx.x * blockDim.x + threadIdx.x;
    if (taskIdx >= p.w * p.h) return;
    const int pixel = __ldg(p.indexToPixel + taskIdx);
    float ns[4];
    ns[0] = 2.0f * ((float)(pixel % p.w) + 0.5f) / (float)p.w - 1.0f;
    ns[1] = 2.0f * ((float)(pixel / p.w) + 0.5f) / (float)p.h - 1.0f;
    ns[2] = 0.0f;
    ns[3] = 1.0f;
    float wp[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        float rr = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; j++) rr += p.m[i + 4 * j] * ns[j];
        wp[i] = rr;
    }
    float dx = wp[0] / wp[3] - p.origin[0];
    float dy = wp[1] / wp[3] - p.origin[1];
    float dz = wp[2] / wp[3] - p.origin[2];
    normalize3(dx, dy, dz);
    p.slotToID[taskIdx] = pixel;
    p.idToSlot[pixel] = taskIdx;
    p.rays[2 * (size_t)taskIdx] = make_float4(p.origin[0], p.origin[1], p.origin[2], 0.0f);
    p.rays[2 * (size_t)taskIdx + 1] = make_float4(dx, dy, dz, p.maxDist);
}

__device__ __forceinline__ void jenkinsMix(uint32_t& a, uint32_t& b, uint32_t& c)
{
    a -= b; a -= c; a ^= (c >> 13);
    b -= c; b -= a; b ^= (a << 8);
    c -= a; c -= b; c ^= (b >> 13);
    a -= b; a -= c; a ^= (c >> 12);
    b -= c; b -= a; b ^= (a << 16);
    c -= a; c -= b; c ^= (b >> 5);
    a -= b; a -= c; a ^= (c >> 3);
    b -= c; b -= a; b ^= (a << 10);
    c -= a; c -= b; c ^= (b >> 15);
}

struct AOParams {
    int firstInputSlot, numInputRays, numSamples;
    float maxDist;
    uint32_t randomSeed;
    const float4* inRays;
    const int4* inResults;
    const float* normals;
    float4* outRays;
    int32_t* outIDToSlot;
    int32_t* outSlotToID;
};

__global__ void __launch_bounds__(256) rayGenAOKernel(const AOParams p)
{
    const long long out = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)p.numInputRays * p.numSamples;
    if (out >= total) return;
    const int taskIdx = (int)(out / p.numSamples);
    const int i = (int)(out - (long long)taskIdx * p.numSamples);
    const int inSlot = taskIdx + p.firstInputSlot;

    const float4 ro = __ldg(p.inRays + 2 * (size_t)inSlot);
    const float4 rd = __ldg(p.inRays + 2 * (size_t)inSlot + 1);
    const int4 res = __ldg(p.inResults + inSlot);

    const float epsilon = 1.0e-4f;
    const float back = fmaxf(__int_as_float(res.y) - epsilon, 0.0f);
    const float ox = ro.x + rd.x * back, oy = ro.y + rd.y * back, oz = ro.z + rd.z * back;

    const int tri = res.x;
    float nx = 1.0f, ny = 0.0f, nz = 0.0f;
    if (tri != -1) { nx = __ldg(p.normals + 3 * (size_t)tri); ny = __ldg(p.normals + 3 * (size_t)tri + 1); nz = __ldg(p.normals + 3 * (size_t)tri + 2); }
    float dn = 0.0f;
    dn += nx * rd.x; dn += ny * rd.y; dn += nz * rd.z;
    if (dn > 0.0f) { nx = -nx; ny = -ny; nz = -nz; }

    const float ax = fabsf(nx), ay = fabsf(ny), az = fabsf(nz);
    const float nm = fmaxf(fmaxf(ax, ay), az);
    float px = ny, py = -nx, pz = 0.0f;                 // assume y is largest
    if (nm == az)      { px = 0.0f; py = nz; pz = -ny; }
    else if (nm == ax) { px = -nz; py = 0.0f; pz = nx; }
    normalize3(px, py, pz);
    const float bx = ny * pz - nz * py, by = nz * px - nx * pz, bz = nx * py - ny * px;   // cross(normal, perp)

    uint32_t hashA = p.randomSeed + (uint32_t)taskIdx, hashB = 0x9e3779b9u, hashC = 0x9e3779b9u;
    jenkinsMix(hashA, hashB, hashC);
    jenkinsMix(hashA, hashB, hashC);
    const float kPi = 3.14159265358979323846f;
    const float angle0 = 2.0f * kPi * (float)hashC * 2.3283064365386963e-10f;
    float sa, ca;
    sincosf(angle0, &sa, &ca);
    const float t0x = px * ca + bx * sa, t0y = py * ca + by * sa, t0z = pz * ca + bz * sa;
    const float t1x = px * -sa + bx * ca, t1y = py * -sa + by * ca, t1z = pz * -sa + bz * ca;

    // Halton base 2 / base 3 of i + 1
    float x = 0.0f, xadd = 1.0f;
    unsigned hc2 = (unsigned)i + 1u;
    while (hc2 != 0) { xadd *= 0.5f; if (hc2 & 1u) x += xadd; hc2 >>= 1; }
    float y = 0.0f, yadd = 1.0f;
    int hc3 = i + 1;
    while (hc3 != 0) { yadd *= 1.0f / 3.0f; y += (float)(hc3 % 3) * yadd; hc3 /= 3; }

    const float angle = 2.0f * kPi * y;
    const float r = sqrtf(x);
    float s1, c1;
    sincosf(angle, &s1, &c1);
    x = r * c1;
    y = r * s1;
    const float z = sqrtf(1.0f - x * x - y * y);

    float dx = x * t0x + y * t1x + z * nx;
    float dy = x * t0y + y * t1y + z * ny;
    float dz = x * t0z + y * t1z + z * nz;
    normalize3(dx, dy, dz);

    p.outRays[2 * (size_t)out] = make_float4(ox, oy, oz, 0.0f);
    p.outRays[2 * (size_t)out + 1] = make_float4(dx, dy, dz, (tri == -1) ? -1.0f : p.maxDist);
    p.outIDToSlot[out] = (int32_t)out;
    p.outSlotToID[out] = (int32_t)out;
}

}  // namespace b2rt

using namespace b2rt;

extern "C" {

int b2rt_raygen_primary(b2rt_ctx* ctx, const float origin[3], const float nscreenToWorld[16], int w, int h,
                        float maxDist, const int32_t* indexToPixel, void* rays, int32_t* idToSlot,
                        int32_t* slotToID, void* stream)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (w < 0 || h < 0 || (long long)w * h > 0x7fffffffLL) return setError(ctx, B2RT_ERR_ARG, "b2rt_raygen_primary: bad size %dx%d", w, h);
    if (w * h == 0) return B2RT_OK;
    if (!origin || !nscreenToWorld || !indexToPixel || !rays || !idToSlot || !slotToID)
        return setError(ctx, B2RT_ERR_ARG, "b2rt_raygen_primary: NULL argument");
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    PrimaryParams p;
    memcpy(p.origin, origin, 12);
    memcpy(p.m, nscreenToWorld, 64);
    p.w = w; p.h = h; p.maxDist = maxDist;
    p.indexToPixel = indexToPixel; p.rays = static_cast<float4*>(rays);
    p.idToSlot = idToSlot; p.slotToID = slotToID;
    const int n = w * h;
    rayGenPrimaryKernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
    B2RT_CHECK_LAUNCH(ctx, "rayGenPrimaryKernel");
    return B2RT_OK;
}

int b2rt_raygen_ao(b2rt_ctx* ctx, int firstInputSlot, int numInputRays, int numSamples, float maxDist,
                   uint32_t randomSeed, const void* inRays, const void* inResults, const float* normals,
                   void* outRays, int32_t* outIDToSlot, int32_t* outSlotToID, void* stream)
{
    if (!ctx) return B2RT_ERR_ARG;
    if (numInputRays < 0 || numSamples < 0 || firstInputSlot < 0 || (long long)numInputRays * numSamples > 0x7fffffffLL)
        return setError(ctx, B2RT_ERR_ARG, "b2rt_raygen_ao: bad counts (%d rays x %d samples)", numInputRays, numSamples);
    const long long total = (long long)numInputRays * numSamples;
    if (total == 0) return B2RT_OK;
    if (!inRays || !inResults || !normals || !outRays || !outIDToSlot || !outSlotToID)
        return setError(ctx, B2RT_ERR_ARG, "b2rt_raygen_ao: NULL argument");
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    AOParams p;
    p.firstInputSlot = firstInputSlot; p.numInputRays = numInputRays; p.numSamples = numSamples;
    p.maxDist = maxDist; p.randomSeed = randomSeed;
    p.inRays = static_cast<const float4*>(inRays); p.inResults = static_cast<const int4*>(inResults);
    p.normals = normals; p.outRays = static_cast<float4*>(outRays);
    p.outIDToSlot = outIDToSlot; p.outSlotToID = outSlotToID;
    rayGenAOKernel<<<(unsigned)((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
    B2RT_CHECK_LAUNCH(ctx, "rayGenAOKernel");
    return B2RT_OK;
}

}  // extern "C"
