/*
 * b2rt_oracle.c -- CPU restatement of the reference hot path (see b2rt_oracle.h).
 * TEST INFRASTRUCTURE ONLY; never linked into the product library.
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared b2rt_oracle.c -o libb2rt_oracle.so -lm
 */
#include "b2rt_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define B2O_F32_MAX 3.402823466e+38f
#define B2O_PI      3.14159265358979323846f   /* FW_PI, src/framework/base/Math.hh */

/* ------------------------------------------------------------------------------------------
 * small vector helpers following src/framework/base/Math.hh (component loops in index order)
 * ---------------------------------------------------------------------------------------- */
typedef struct { float x, y, z; } v3;

static inline v3 v3make(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 v3add(v3 a, v3 b) { return v3make(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 v3sub(v3 a, v3 b) { return v3make(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 v3scale(v3 a, float s) { return v3make(a.x * s, a.y * s, a.z * s); }
static inline v3 v3neg(v3 a) { return v3make(-a.x, -a.y, -a.z); }
/* Math.hh:140 lenSqr accumulates r += sqr(tp[i]) from zero; dot does the same with products */
static inline float v3dot(v3 a, v3 b) { float r = 0.0f; r += a.x * b.x; r += a.y * b.y; r += a.z * b.z; return r; }
static inline v3 v3cross(v3 a, v3 b)
{
    return v3make(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
/* Math.hh:113 rcp(a) = a ? 1/a : 0 ;  Math.hh:142 normalized = v * (1 * rcp(length)) */
static inline float frcp(float a) { return (a != 0.0f) ? 1.0f / a : 0.0f; }
static inline v3 v3normalize(v3 a) { return v3scale(a, 1.0f * frcp(sqrtf(v3dot(a, a)))); }
/* Defs.hh:160 host min/max: (a < b) ? a : b  /  (a > b) ? a : b */
static inline float fwmin(float a, float b) { return (a < b) ? a : b; }
static inline float fwmax(float a, float b) { return (a > b) ? a : b; }

/* ------------------------------------------------------------------------------------------
 * PixelTable::recalculate -- src/rt/ray/PixelTable.cc:94-156
 * 8x8 blocks visited in Morton order, Morton order inside a block, then the bottom stripe,
 * then the right stripe with the corner.
 * ---------------------------------------------------------------------------------------- */
void b2o_pixel_table(int w, int h, int32_t* indexToPixel, int32_t* pixelToIndex)
{
    int idx = 0;
    int bheight = h & ~7;
    int bwidth = w & ~7;
    int maxdim = (bwidth > bheight) ? bwidth : bheight;
    maxdim |= maxdim >> 1;
    maxdim |= maxdim >> 2;
    maxdim |= maxdim >> 4;
    maxdim |= maxdim >> 8;
    maxdim |= maxdim >> 16;
    maxdim = (maxdim + 1) >> 1;

    int width8 = bwidth >> 3;
    int height8 = bheight >> 3;
    for (long long i = 0; i < (long long)maxdim * maxdim; i++) {
        int tx = 0, ty = 0, bit = 1;
        long long val = i;
        while (val) {
            if (val & 1) tx |= bit;
            if (val & 2) ty |= bit;
            bit += bit;
            val >>= 2;
        }
        if (tx < width8 && ty < height8) {
            for (int inner = 0; inner < 64; inner++) {
                int ix = ((inner & 1) >> 0) | ((inner & 4) >> 1) | ((inner & 16) >> 2);
                int iy = ((inner & 2) >> 1) | ((inner & 8) >> 2) | ((inner & 32) >> 3);
                int pos = (ty * 8 + iy) * w + (tx * 8 + ix);
                if (pixelToIndex) pixelToIndex[pos] = idx;
                indexToPixel[idx++] = pos;
            }
        }
    }
    for (int px = 0; px < bwidth; px++)
        for (int py = bheight; py < h; py++) {
            int pos = px + py * w;
            if (pixelToIndex) pixelToIndex[pos] = idx;
            indexToPixel[idx++] = pos;
        }
    for (int py = 0; py < h; py++)
        for (int px = bwidth; px < w; px++) {
            int pos = px + py * w;
            if (pixelToIndex) pixelToIndex[pos] = idx;
            indexToPixel[idx++] = pos;
        }
}

/* ------------------------------------------------------------------------------------------
 * 4x4 matrices, column-major like FW::Mat4f (Math.hh:520: get(r,c) = m[r + 4c]).
 * ---------------------------------------------------------------------------------------- */
#define M(m, r, c) ((m)[(r) + 4 * (c)])

static void m4identity(float* m) { memset(m, 0, 16 * sizeof(float)); M(m,0,0) = M(m,1,1) = M(m,2,2) = M(m,3,3) = 1.0f; }

/* Math.hh:1034-1048 */
static void m4mul(const float* a, const float* b, float* out)
{
    float r[16];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float rr = 0.0f;
            for (int k = 0; k < 4; k++) rr += M(a, i, k) * M(b, k, j);
            M(r, i, j) = rr;
        }
    memcpy(out, r, sizeof(r));
}

/* Math.hh:935-940 */
static float det3(const float s[9]) /* s column-major 3x3: s[r + 3c] */
{
#define S3(r, c) s[(r) + 3 * (c)]
    return S3(0,0) * S3(1,1) * S3(2,2) - S3(0,0) * S3(1,2) * S3(2,1) +
           S3(1,0) * S3(2,1) * S3(0,2) - S3(1,0) * S3(2,2) * S3(0,1) +
           S3(2,0) * S3(0,1) * S3(1,2) - S3(2,0) * S3(0,2) * S3(1,1);
#undef S3
}

/* MatrixBase::inverted -- Math.hh:962-983.  Cofactor expansion; `d` accumulates L copies of
 * the determinant, hence the trailing `* L`. */
static void m4invert(const float* m, float* out)
{
    float r[16];
    float d = 0.0f;
    float si = 1.0f;
    for (int i = 0; i < 4; i++) {
        float sj = si;
        for (int j = 0; j < 4; j++) {
            float sub[9];
            for (int k = 0; k < 3; k++)
                for (int l = 0; l < 3; l++)
                    sub[k + 3 * l] = M(m, (k < j) ? k : k + 1, (l < i) ? l : l + 1);
            float dd = det3(sub) * sj;
            M(r, i, j) = dd;
            d += dd * M(m, j, i);
            sj = -sj;
        }
        si = -si;
    }
    float rd = frcp(d);
    for (int k = 0; k < 16; k++) out[k] = r[k] * rd * 4.0f;
}

/* CameraControls::getOrientation/getWorldToCamera -- CameraControls.cc:263-297;
 * Mat4f::perspective -- Math.cc:78-91; Mat4f::fitToView -- Math.cc:64-75;
 * Renderer::beginFrame -- Renderer.cc:126-129: nscreenToWorld =
 *     invert(fitToView(-1, 2, viewSize) * (perspective * worldToCamera)) */
void b2o_camera_matrices(const float pos[3], const float forward[3], const float up[3],
                         float fov, float znear, float zfar, int w, int h,
                         float origin[3], float nscreenToWorld[16])
{
    v3 p = v3make(pos[0], pos[1], pos[2]);
    v3 c2 = v3neg(v3normalize(v3make(forward[0], forward[1], forward[2])));
    v3 c0 = v3normalize(v3cross(v3make(up[0], up[1], up[2]), c2));
    v3 c1 = v3normalize(v3cross(c2, c0));
    /* pos' = orient^T * position (Math.hh:994-1006: rr += get(i,j) * v[j]) */
    float px = 0.0f, py = 0.0f, pz = 0.0f;
    px += c0.x * p.x; px += c0.y * p.y; px += c0.z * p.z;
    py += c1.x * p.x; py += c1.y * p.y; py += c1.z * p.z;
    pz += c2.x * p.x; pz += c2.y * p.y; pz += c2.z * p.z;

    float w2c[16]; m4identity(w2c);
    M(w2c,0,0) = c0.x; M(w2c,0,1) = c0.y; M(w2c,0,2) = c0.z; M(w2c,0,3) = -px;
    M(w2c,1,0) = c1.x; M(w2c,1,1) = c1.y; M(w2c,1,2) = c1.z; M(w2c,1,3) = -py;
    M(w2c,2,0) = c2.x; M(w2c,2,1) = c2.y; M(w2c,2,2) = c2.z; M(w2c,2,3) = -pz;

    float f = frcp(tanf(fov * B2O_PI / 360.0f));
    float d = frcp(znear - zfar);
    float persp[16]; memset(persp, 0, sizeof(persp));
    M(persp,0,0) = f; M(persp,1,1) = f;
    M(persp,2,2) = (znear + zfar) * d; M(persp,2,3) = 2.0f * znear * zfar * d;
    M(persp,3,2) = -1.0f;

    float w2clip[16]; m4mul(persp, w2c, w2clip);

    /* fitToView(pos=-1, size=2, viewSize=(w,h)) = scale(2/viewSize,1) * scale(min(viewSize/size),1)
     *                                             * translate(-pos - size/2, 0) */
    float vw = (float)w, vh = (float)h;
    float s1[16]; m4identity(s1); M(s1,0,0) = 2.0f / vw; M(s1,1,1) = 2.0f / vh; M(s1,2,2) = 1.0f;
    float mn = fwmin(vw / 2.0f, vh / 2.0f);
    float s2[16]; m4identity(s2); M(s2,0,0) = mn; M(s2,1,1) = mn; M(s2,2,2) = 1.0f;
    float tr[16]; m4identity(tr);
    M(tr,0,3) = -(-1.0f) - 2.0f * 0.5f; M(tr,1,3) = -(-1.0f) - 2.0f * 0.5f; M(tr,2,3) = 0.0f;
    float fit[16], tmp[16];
    m4mul(s1, s2, tmp); m4mul(tmp, tr, fit);
    m4mul(fit, w2clip, tmp);
    m4invert(tmp, nscreenToWorld);
    origin[0] = pos[0]; origin[1] = pos[1]; origin[2] = pos[2];
}

/* ------------------------------------------------------------------------------------------
 * rayGenPrimaryKernel -- src/rt/ray/RayGenKernels.cu:79-113
 * ---------------------------------------------------------------------------------------- */
void b2o_raygen_primary(const float origin[3], const float nscreenToWorld[16], int w, int h,
                        float maxDist, const int32_t* indexToPixel,
                        b2o_ray* rays, int32_t* idToSlot, int32_t* slotToID)
{
    const float* m = nscreenToWorld;
    v3 org = v3make(origin[0], origin[1], origin[2]);
    for (int taskIdx = 0; taskIdx < w * h; taskIdx++) {
        int pixel = indexToPixel[taskIdx];
        float ns[4];
        ns[0] = 2.0f * ((float)(pixel % w) + 0.5f) / (float)w - 1.0f;
        ns[1] = 2.0f * ((float)(pixel / w) + 0.5f) / (float)h - 1.0f;
        ns[2] = 0.0f;
        ns[3] = 1.0f;
        float wp[4];
        for (int i = 0; i < 4; i++) {      /* Math.hh:990-1006 */
            float rr = 0.0f;
            for (int j = 0; j < 4; j++) rr += M(m, i, j) * ns[j];
            wp[i] = rr;
        }
        v3 worldPos = v3make(wp[0] / wp[3], wp[1] / wp[3], wp[2] / wp[3]);
        v3 dir = v3normalize(v3sub(worldPos, org));
        slotToID[taskIdx] = pixel;
        idToSlot[pixel] = taskIdx;
        b2o_ray* r = &rays[taskIdx];
        r->ox = org.x; r->oy = org.y; r->oz = org.z; r->tmin = 0.0f;
        r->dx = dir.x; r->dy = dir.y; r->dz = dir.z; r->tmax = maxDist;
    }
}

/* jenkinsMix -- RayGenKernels.cu:36-47 */
static inline void jenkins_mix(uint32_t* pa, uint32_t* pb, uint32_t* pc)
{
    uint32_t a = *pa, b = *pb, c = *pc;
    a -= b; a -= c; a ^= (c >> 13);
    b -= c; b -= a; b ^= (a << 8);
    c -= a; c -= b; c ^= (b >> 13);
    a -= b; a -= c; a ^= (c >> 12);
    b -= c; b -= a; b ^= (a << 16);
    c -= a; c -= b; c ^= (b >> 5);
    a -= b; a -= c; a ^= (c >> 3);
    b -= c; b -= a; b ^= (a << 10);
    c -= a; c -= b; c ^= (b >> 15);
    *pa = a; *pb = b; *pc = c;
}

/* ------------------------------------------------------------------------------------------
 * rayGenAOKernel -- src/rt/ray/RayGenKernels.cu:117-227 (AO and Diffuse share it,
 * Renderer.cc:268-277).  libm sinf/cosf stand in for the device's fast-math intrinsics.
 * ---------------------------------------------------------------------------------------- */
void b2o_raygen_ao(int firstInputSlot, int numInputRays, int numSamples, float maxDist,
                   uint32_t randomSeed, const b2o_ray* inRays, const b2o_result* inResults,
                   const float* normals, b2o_ray* outRaysBase, int32_t* outIDToSlotBase,
                   int32_t* outSlotToIDBase)
{
    for (int taskIdx = 0; taskIdx < numInputRays; taskIdx++) {
        int inSlot = taskIdx + firstInputSlot;
        const b2o_ray* inRay = &inRays[inSlot];
        const b2o_result* inResult = &inResults[inSlot];
        int outSlot = taskIdx * numSamples;
        b2o_ray* outRays = outRaysBase + outSlot;
        int32_t* outIDToSlot = outIDToSlotBase + outSlot;
        int32_t* outSlotToID = outSlotToIDBase + outSlot;

        const float epsilon = 1.0e-4f;
        v3 io = v3make(inRay->ox, inRay->oy, inRay->oz);
        v3 id = v3make(inRay->dx, inRay->dy, inRay->dz);
        v3 origin = v3add(io, v3scale(id, fmaxf(inResult->t - epsilon, 0.0f)));

        int tri = inResult->id;
        v3 normal = v3make(1.0f, 0.0f, 0.0f);
        if (tri != -1) normal = v3make(normals[3 * tri], normals[3 * tri + 1], normals[3 * tri + 2]);
        if (v3dot(normal, id) > 0.0f) normal = v3neg(normal);

        v3 na = v3make(fabsf(normal.x), fabsf(normal.y), fabsf(normal.z));
        float nm = fmaxf(fmaxf(na.x, na.y), na.z);
        v3 perp = v3make(normal.y, -normal.x, 0.0f);
        if (nm == na.z)      perp = v3make(0.0f, normal.z, -normal.y);
        else if (nm == na.x) perp = v3make(-normal.z, 0.0f, normal.x);
        perp = v3normalize(perp);
        v3 biperp = v3cross(normal, perp);

        uint32_t hashA = randomSeed + (uint32_t)taskIdx;
        uint32_t hashB = 0x9e3779b9u;
        uint32_t hashC = 0x9e3779b9u;
        jenkins_mix(&hashA, &hashB, &hashC);
        jenkins_mix(&hashA, &hashB, &hashC);
        float angle = 2.0f * B2O_PI * (float)hashC * 2.3283064365386963e-10f; /* exp2(-32) */

        float ca = cosf(angle), sa = sinf(angle);
        v3 t0 = v3add(v3scale(perp, ca), v3scale(biperp, sa));
        v3 t1 = v3add(v3scale(perp, -sa), v3scale(biperp, ca));

        for (int i = 0; i < numSamples; i++) {
            float x = 0.0f, xadd = 1.0f;
            unsigned int hc2 = (unsigned int)i + 1;
            while (hc2 != 0) {
                xadd *= 0.5f;
                if ((hc2 & 1) != 0) x += xadd;
                hc2 >>= 1;
            }
            float y = 0.0f, yadd = 1.0f;
            int hc3 = i + 1;
            while (hc3 != 0) {
                yadd *= 1.0f / 3.0f;
                y += (float)(hc3 % 3) * yadd;
                hc3 /= 3;
            }
            float ang = 2.0f * B2O_PI * y;
            float r = sqrtf(x);
            x = r * cosf(ang);
            y = r * sinf(ang);
            float z = sqrtf(1.0f - x * x - y * y);

            v3 dir = v3normalize(v3add(v3add(v3scale(t0, x), v3scale(t1, y)), v3scale(normal, z)));
            outRays[i].ox = origin.x; outRays[i].oy = origin.y; outRays[i].oz = origin.z;
            outRays[i].dx = dir.x; outRays[i].dy = dir.y; outRays[i].dz = dir.z;
            outRays[i].tmin = 0.0f;
            outRays[i].tmax = (tri == -1) ? -1.0f : maxDist;
            outIDToSlot[i] = i + outSlot;
            outSlotToID[i] = i + outSlot;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * findAABBKernel -- src/rt/ray/RayBufferKernels.cu:74-140.  min/max are order independent
 * for non-NaN input, so a serial loop gives the kernel's result.  Device min/max on F32 are
 * fminf/fmaxf (Defs.hh:166).
 * ---------------------------------------------------------------------------------------- */
void b2o_find_aabb(const b2o_ray* rays, int n, float lo[3], float hi[3])
{
    lo[0] = lo[1] = lo[2] = +B2O_F32_MAX;
    hi[0] = hi[1] = hi[2] = -B2O_F32_MAX;
    for (int i = 0; i < n; i++) {
        const b2o_ray* r = &rays[i];
        float o[3] = { r->ox, r->oy, r->oz };
        float d[3] = { r->dx, r->dy, r->dz };
        for (int k = 0; k < 3; k++) {
            lo[k] = fminf(lo[k], o[k]); hi[k] = fmaxf(hi[k], o[k]);
            float e = o[k] + d[k] * r->tmax;
            lo[k] = fminf(lo[k], e); hi[k] = fmaxf(hi[k], e);
        }
    }
}

/* collectBits -- RayBufferKernels.cu:66-70: bit i of `value` lands at key bit index + 6 i */
static inline void collect_bits(uint32_t* hash, int index, uint32_t value)
{
    for (int i = 0; i < 32; i++)
        hash[(index + i * 6) >> 5] |= ((value >> i) & 1u) << ((index + i * 6) & 31);
}

/* float -> U32 as the device does it (cvt.rzi.u32.f32 saturates: negative/NaN -> 0,
 * too large -> 0xffffffff); C's cast is undefined out of range, so spell it out. */
static inline uint32_t f2u_sat(float f)
{
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967296.0f) return 0xffffffffu;
    return (uint32_t)f;
}

/* genMortonKeysKernel -- RayBufferKernels.cu:144-179 */
void b2o_morton_keys(const b2o_ray* rays, int n, const float lo[3], const float hi[3],
                     b2o_mortonkey* keys)
{
    for (int i = 0; i < n; i++) {
        const b2o_ray* r = &rays[i];
        float a[3], b[3];
        a[0] = (r->ox - lo[0]) / (hi[0] - lo[0]);
        a[1] = (r->oy - lo[1]) / (hi[1] - lo[1]);
        a[2] = (r->oz - lo[2]) / (hi[2] - lo[2]);
        v3 nd = v3normalize(v3make(r->dx, r->dy, r->dz));
        b[0] = (nd.x + 1.0f) * 0.5f;
        b[1] = (nd.y + 1.0f) * 0.5f;
        b[2] = (nd.z + 1.0f) * 0.5f;
        b2o_mortonkey* k = &keys[i];
        k->oldSlot = i;
        for (int j = 0; j < 6; j++) k->hash[j] = 0;
        collect_bits(k->hash, 0, f2u_sat(a[0] * 256.0f * 65536.0f));
        collect_bits(k->hash, 1, f2u_sat(a[1] * 256.0f * 65536.0f));
        collect_bits(k->hash, 2, f2u_sat(a[2] * 256.0f * 65536.0f));
        collect_bits(k->hash, 3, f2u_sat(b[0] * 32.0f * 65536.0f));
        collect_bits(k->hash, 4, f2u_sat(b[1] * 32.0f * 65536.0f));
        collect_bits(k->hash, 5, f2u_sat(b[2] * 32.0f * 65536.0f));
    }
}

/* compareMortonKey -- RayBuffer.cc:237-248: most significant word first */
int b2o_key_less(const b2o_mortonkey* a, const b2o_mortonkey* b)
{
    for (int j = 5; j >= 0; j--)
        if (a->hash[j] != b->hash[j]) return a->hash[j] < b->hash[j];
    return 0;
}

/* The reference sorts with an unstable quicksort (Sort.cc:131-165), so the order inside a run
 * of equal keys is unspecified.  The oracle uses a stable merge sort: its output is ONE of the
 * orders the reference may produce; parity tests compare key sequences, not slot ties. */
static void merge_sort_keys(b2o_mortonkey* a, b2o_mortonkey* tmp, int n)
{
    if (n < 2) return;
    int h = n / 2;
    merge_sort_keys(a, tmp, h);
    merge_sort_keys(a + h, tmp, n - h);
    int i = 0, j = h, k = 0;
    while (i < h && j < n) tmp[k++] = b2o_key_less(&a[j], &a[i]) ? a[j++] : a[i++];
    while (i < h) tmp[k++] = a[i++];
    while (j < n) tmp[k++] = a[j++];
    memcpy(a, tmp, (size_t)n * sizeof(b2o_mortonkey));
}

void b2o_sort_keys(b2o_mortonkey* keys, int n)
{
    b2o_mortonkey* tmp = (b2o_mortonkey*)malloc((size_t)(n > 0 ? n : 1) * sizeof(b2o_mortonkey));
    merge_sort_keys(keys, tmp, n);
    free(tmp);
}

/* reorderRaysKernel -- RayBufferKernels.cu:183-200 */
void b2o_reorder_rays(int n, const b2o_mortonkey* keys, const b2o_ray* inRays,
                      const int32_t* inSlotToID, b2o_ray* outRays, int32_t* outIDToSlot,
                      int32_t* outSlotToID)
{
    for (int i = 0; i < n; i++) {
        int oldSlot = keys[i].oldSlot;
        int id = inSlotToID[oldSlot];
        outRays[i] = inRays[oldSlot];
        outIDToSlot[id] = i;
        outSlotToID[i] = id;
    }
}

/* ------------------------------------------------------------------------------------------
 * CudaBVH::woopifyTri -- src/rt/cuda/CudaBVH.cc:361-380, Mat4f inverse Math.hh:962-983.
 * mtx columns: (v0-v2,0), (v1-v2,0), (cross(v0-v2, v1-v2),0), (v2,1).
 * woop[0..3]  = (inv(2,0), inv(2,1), inv(2,2), -inv(2,3))
 * woop[4..7]  = row 0 of inv,  woop[8..11] = row 1 of inv.
 * CudaBVH.cc:316-317 rewrites -0.0f in woop[0] as +0.0f to protect the leaf terminator.
 * ---------------------------------------------------------------------------------------- */
void b2o_woopify(const float p0[3], const float p1[3], const float p2[3], float woop[12])
{
    v3 v0 = v3make(p0[0], p0[1], p0[2]);
    v3 v1 = v3make(p1[0], p1[1], p1[2]);
    v3 v2 = v3make(p2[0], p2[1], p2[2]);
    v3 e0 = v3sub(v0, v2), e1 = v3sub(v1, v2);
    v3 n = v3cross(e0, e1);
    float m[16], inv[16];
    M(m,0,0) = e0.x; M(m,1,0) = e0.y; M(m,2,0) = e0.z; M(m,3,0) = 0.0f;
    M(m,0,1) = e1.x; M(m,1,1) = e1.y; M(m,2,1) = e1.z; M(m,3,1) = 0.0f;
    M(m,0,2) = n.x;  M(m,1,2) = n.y;  M(m,2,2) = n.z;  M(m,3,2) = 0.0f;
    M(m,0,3) = v2.x; M(m,1,3) = v2.y; M(m,2,3) = v2.z; M(m,3,3) = 1.0f;
    m4invert(m, inv);
    woop[0] = M(inv,2,0); woop[1] = M(inv,2,1); woop[2] = M(inv,2,2); woop[3] = -M(inv,2,3);
    for (int c = 0; c < 4; c++) { woop[4 + c] = M(inv,0,c); woop[8 + c] = M(inv,1,c); }
    if (woop[0] == 0.0f) woop[0] = 0.0f;
}

/* Scene::Scene -- src/rt/Scene.cc:72: normalize(cross(v1 - v0, v2 - v0)) */
void b2o_tri_normals(const float* verts, const int32_t* tris, int numTris, float* normals)
{
    for (int i = 0; i < numTris; i++) {
        const float* a = verts + 3 * tris[3 * i];
        const float* b = verts + 3 * tris[3 * i + 1];
        const float* c = verts + 3 * tris[3 * i + 2];
        v3 va = v3make(a[0], a[1], a[2]);
        v3 nrm = v3normalize(v3cross(v3sub(v3make(b[0], b[1], b[2]), va), v3sub(v3make(c[0], c[1], c[2]), va)));
        normals[3 * i] = nrm.x; normals[3 * i + 1] = nrm.y; normals[3 * i + 2] = nrm.z;
    }
}

/* ------------------------------------------------------------------------------------------
 * Intersect::RayBox -- src/rt/Util.cc:34-46 (true division, FW host min/max semantics)
 * ---------------------------------------------------------------------------------------- */
static inline void ray_box(const float lo[3], const float hi[3], const float o[3], const float d[3],
                           float* tmin, float* tmax)
{
    float mn[3], mx[3];
    for (int k = 0; k < 3; k++) {
        float t0 = (lo[k] - o[k]) / d[k];
        float t1 = (hi[k] - o[k]) / d[k];
        mn[k] = fwmin(t0, t1);
        mx[k] = fwmax(t0, t1);
    }
    float a = mn[0]; a = fwmax(a, mn[1]); a = fwmax(a, mn[2]);   /* Math.hh:145 */
    float b = mx[0]; b = fwmin(b, mx[1]); b = fwmin(b, mx[2]);   /* Math.hh:144 */
    *tmin = a; *tmax = b;
}

/* Intersect::RayTriangle -- src/rt/Util.cc:50-94 (Moller-Trumbore, EPSILON = 0, two sided).
 * Returns t (or F32_MAX on a miss); u,v through pointers. */
static inline float ray_triangle_mt(v3 v0, v3 v1, v3 v2, v3 org, v3 dir, float rtmin, float rtmax,
                                    float* pu, float* pv)
{
    v3 edge1 = v3sub(v1, v0);
    v3 edge2 = v3sub(v2, v0);
    v3 pvec = v3cross(dir, edge2);
    float det = v3dot(edge1, pvec);
    v3 tvec = v3sub(org, v0);
    float u = v3dot(tvec, pvec);
    v3 qvec = v3cross(tvec, edge1);
    float v = v3dot(dir, qvec);
    if (det > 0.0f) {
        if (u < 0.0 || u > det) return B2O_F32_MAX;
        if (v < 0.0 || u + v > det) return B2O_F32_MAX;
    } else if (det < -0.0f) {
        if (u > 0.0 || u < det) return B2O_F32_MAX;
        if (v > 0.0 || u + v < det) return B2O_F32_MAX;
    } else
        return B2O_F32_MAX;
    float inv_det = 1.f / det;
    float t = v3dot(edge2, qvec) * inv_det;
    u *= inv_det;
    v *= inv_det;
    if (t > rtmin && t < rtmax) { *pu = u; *pv = v; return t; }
    return B2O_F32_MAX;
}

typedef struct {
    const float* nodes; const float* woop; const int32_t* triIndex;
    const float* verts; const int32_t* tris;
    v3 org, dir; float o[3], d[3]; float tmin, tmax;
    b2o_result* result; int needClosestHit;
} mt_ctx;

/* BVH::traceRecursive -- src/rt/bvh/BVH.cc:89-163.  `addr` is a Compact2 child reference:
 * >= 0 inner node (float4 index), < 0 leaf (~float4 index of its first Woop row). */
static void mt_recurse(mt_ctx* c, int addr)
{
    if (addr < 0) {
        for (int row = ~addr;; row += 3) {
            uint32_t bits; memcpy(&bits, &c->woop[4 * row], 4);
            if (bits == 0x80000000u) break;                       /* leaf terminator */
            int index = c->triIndex[row];
            const int32_t* ind = &c->tris[3 * index];
            const float* a = &c->verts[3 * ind[0]];
            const float* b = &c->verts[3 * ind[1]];
            const float* cc = &c->verts[3 * ind[2]];
            float u = 0, v = 0;
            float t = ray_triangle_mt(v3make(a[0], a[1], a[2]), v3make(b[0], b[1], b[2]),
                                      v3make(cc[0], cc[1], cc[2]), c->org, c->dir, c->tmin, c->tmax, &u, &v);
            if (t > c->tmin && t < c->tmax) {
                c->tmax = t;
                c->result->t = t;
                c->result->id = index;
                if (!c->needClosestHit) return;
            }
        }
        return;
    }
    const float* n = &c->nodes[4 * addr];
    int32_t child[2]; memcpy(child, &n[12], 8);
    float lo0[3] = { n[0], n[2], n[8] },  hi0[3] = { n[1], n[3], n[9] };
    float lo1[3] = { n[4], n[6], n[10] }, hi1[3] = { n[5], n[7], n[11] };
    float s0min, s0max, s1min, s1max;
    ray_box(lo0, hi0, c->o, c->d, &s0min, &s0max);
    ray_box(lo1, hi1, c->o, c->d, &s1min, &s1max);
    int i0 = (s0min <= s0max) && (s0max >= c->tmin) && (s0min <= c->tmax);
    int i1 = (s1min <= s1max) && (s1max >= c->tmin) && (s1min <= c->tmax);
    int c0 = child[0], c1 = child[1];
    if (i0 && i1)
        if (s0min > s1min) { int t = c0; c0 = c1; c1 = t; }
    if (i0) mt_recurse(c, c0);
    if (c->result->id != -1 && !c->needClosestHit) return;
    if (i1) mt_recurse(c, c1);
}

/* BVH::trace -- src/rt/bvh/BVH.cc:67-87 */
void b2o_trace_mt(const float* nodes, const float* woop, const int32_t* triIndex, const float* verts,
                  const int32_t* tris, const b2o_ray* rays, b2o_result* results, int n,
                  int needClosestHit)
{
    for (int i = 0; i < n; i++) {
        mt_ctx c;
        c.nodes = nodes; c.woop = woop; c.triIndex = triIndex; c.verts = verts; c.tris = tris;
        c.org = v3make(rays[i].ox, rays[i].oy, rays[i].oz);
        c.dir = v3make(rays[i].dx, rays[i].dy, rays[i].dz);
        c.o[0] = rays[i].ox; c.o[1] = rays[i].oy; c.o[2] = rays[i].oz;
        c.d[0] = rays[i].dx; c.d[1] = rays[i].dy; c.d[2] = rays[i].dz;
        c.tmin = rays[i].tmin; c.tmax = rays[i].tmax;
        c.result = &results[i]; c.needClosestHit = needClosestHit;
        results[i].id = -1;                                       /* RayResult::clear */
        mt_recurse(&c, 0);
    }
}

/* ------------------------------------------------------------------------------------------
 * Scalar restatement of the GPU tracer -- src/rt/kernels/kepler_dynamic_fetch.cu:
 * ray setup :123-140, node test :213-285, Woop test :328-381, result :407-408.
 * Canonical order: no speculation, no postponement (neither changes the result of a
 * closest-hit query).  fmaf marks the products nvcc contracts (-fmad=true, the reference build); in the
 * three direction sums dx*x + dy*y + dz*z the plain product is the y term, as in the SASS of the reference
 * kernel built here (oracle/_ref/libref_gpu.so: FMUL dy*y, FFMA dx*x, FFMA dz*z -- fast-math re-association).
 * spanBegin/EndKepler (CudaTracerKernels.hh:274-275) are decision-identical to fminf/fmaxf
 * for tmin >= 0; plain float min/max is used here.
 * ---------------------------------------------------------------------------------------- */
void b2o_trace_woop(const float* nodes, const float* woop, const int32_t* triIndex,
                    const b2o_ray* rays, b2o_result* results, int n, int anyHit,
                    b2o_counters* ctr)
{
    int stack[128];
    for (int i = 0; i < n; i++) {
        const b2o_ray* r = &rays[i];
        float origx = r->ox, origy = r->oy, origz = r->oz, tmin = r->tmin;
        float dirx = r->dx, diry = r->dy, dirz = r->dz, hitT = r->tmax;
        float ooeps = exp2f(-80.0f);
        float idirx = 1.0f / (fabsf(dirx) > ooeps ? dirx : copysignf(ooeps, dirx));
        float idiry = 1.0f / (fabsf(diry) > ooeps ? diry : copysignf(ooeps, diry));
        float idirz = 1.0f / (fabsf(dirz) > ooeps ? dirz : copysignf(ooeps, dirz));
        float oodx = origx * idirx, oody = origy * idiry, oodz = origz * idirz;

        int sp = 0;
        stack[0] = B2O_SENTINEL;
        int nodeAddr = 0, hitIndex = -1;
        float hitU = 0.0f, hitV = 0.0f;
        if (ctr) ctr->rays++;

        while (nodeAddr != B2O_SENTINEL) {
            if (nodeAddr >= 0) {
                const float* nd = &nodes[4 * nodeAddr];
                int32_t cn[2]; memcpy(cn, &nd[12], 8);
                if (ctr) ctr->nodes++;
                float c0lox = fmaf(nd[0], idirx, -oodx), c0hix = fmaf(nd[1], idirx, -oodx);
                float c0loy = fmaf(nd[2], idiry, -oody), c0hiy = fmaf(nd[3], idiry, -oody);
                float c0loz = fmaf(nd[8], idirz, -oodz), c0hiz = fmaf(nd[9], idirz, -oodz);
                float c1loz = fmaf(nd[10], idirz, -oodz), c1hiz = fmaf(nd[11], idirz, -oodz);
                float c1lox = fmaf(nd[4], idirx, -oodx), c1hix = fmaf(nd[5], idirx, -oodx);
                float c1loy = fmaf(nd[6], idiry, -oody), c1hiy = fmaf(nd[7], idiry, -oody);
                float c0min = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), tmin));
                float c0max = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), hitT));
                float c1min = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), tmin));
                float c1max = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), hitT));
                int swp = (c1min < c0min);
                int t0 = (c0max >= c0min), t1 = (c1max >= c1min);
                if (!t0 && !t1) {
                    nodeAddr = stack[sp--];
                } else {
                    nodeAddr = t0 ? cn[0] : cn[1];
                    if (t0 && t1) {
                        int far = cn[1];
                        if (swp) { far = nodeAddr; nodeAddr = cn[1]; }
                        stack[++sp] = far;
                        if (ctr && sp > ctr->maxStack) ctr->maxStack = sp;
                    }
                }
            } else {
                if (ctr) ctr->leaves++;
                int done = 0;
                for (int triAddr = ~nodeAddr;; triAddr += 3) {
                    const float* v00 = &woop[4 * triAddr];
                    uint32_t bits; memcpy(&bits, v00, 4);
                    if (bits == 0x80000000u) break;
                    const float* v11 = v00 + 4;
                    const float* v22 = v00 + 8;
                    if (ctr) ctr->tris++;
                    float Oz = fmaf(-origz, v00[2], fmaf(-origy, v00[1], fmaf(-origx, v00[0], v00[3])));
                    float invDz = 1.0f / fmaf(dirz, v00[2], fmaf(dirx, v00[0], diry * v00[1]));
                    float t = Oz * invDz;
                    if (t > tmin && t < hitT) {
                        if (ctr) ctr->ustage++;
                        float Ox = fmaf(origz, v11[2], fmaf(origy, v11[1], fmaf(origx, v11[0], v11[3])));
                        float Dx = fmaf(dirz, v11[2], fmaf(dirx, v11[0], diry * v11[1]));
                        float u = fmaf(t, Dx, Ox);
                        if (u >= 0.0f) {
                            if (ctr) ctr->vstage++;
                            float Oy = fmaf(origz, v22[2], fmaf(origy, v22[1], fmaf(origx, v22[0], v22[3])));
                            float Dy = fmaf(dirz, v22[2], fmaf(dirx, v22[0], diry * v22[1]));
                            float v = fmaf(t, Dy, Oy);
                            if (v >= 0.0f && u + v <= 1.0f) {
                                hitT = t; hitIndex = triAddr; hitU = u; hitV = v;
                                if (anyHit) { done = 1; break; }
                            }
                        }
                    }
                }
                nodeAddr = done ? B2O_SENTINEL : stack[sp--];
            }
        }
        results[i].id = (hitIndex == -1) ? -1 : triIndex[hitIndex];
        results[i].t = hitT;
        memcpy(&results[i].padA, &hitU, 4);
        memcpy(&results[i].padB, &hitV, 4);
        if (ctr && hitIndex != -1) ctr->hits++;
    }
}

/* countHitsKernel -- src/rt/cuda/RendererKernels.cu:112-162 */
int64_t b2o_count_hits(const b2o_result* results, int64_t n)
{
    int64_t c = 0;
    for (int64_t i = 0; i < n; i++) if (results[i].id >= 0) c++;
    return c;
}

/* Vec4f::fromABGR / toABGR -- src/framework/base/Math.cc:34-54 */
static inline void from_abgr(uint32_t abgr, float c[4])
{
    c[0] = (float)(abgr & 0xFF) * (1.0f / 255.0f);
    c[1] = (float)((abgr >> 8) & 0xFF) * (1.0f / 255.0f);
    c[2] = (float)((abgr >> 16) & 0xFF) * (1.0f / 255.0f);
    c[3] = (float)(abgr >> 24) * (1.0f / 255.0f);
}
static inline uint32_t to_abgr_chan(float x)
{
    float cl = fwmin(fwmax(x, 0.0f), 1.0f);
    return (((uint32_t)(((uint64_t)(cl * 72057594037927936.0f /* 2^56 */) * 255) >> 55) + 1) >> 1);
}
static inline uint32_t to_abgr(const float c[4])
{
    return (to_abgr_chan(c[0]) << 0) | (to_abgr_chan(c[1]) << 8) | (to_abgr_chan(c[2]) << 16) | (to_abgr_chan(c[3]) << 24);
}

/* reconstructKernel -- src/rt/cuda/RendererKernels.cu:60-108 */
void b2o_reconstruct(int numRaysPerPrimary, int firstPrimary, int numPrimary, int isPrimary,
                     int isAO, int isDiffuse, const int32_t* primarySlotToID,
                     const b2o_result* primaryResults, const int32_t* batchIDToSlot,
                     const b2o_result* batchResults, const uint32_t* triMaterialColor,
                     const uint32_t* triShadedColor, uint32_t* pixels)
{
    const float bg[4] = { 0.2f, 0.4f, 0.8f, 1.0f };
    for (int taskIdx = 0; taskIdx < numPrimary; taskIdx++) {
        int primarySlot = firstPrimary + taskIdx;
        int primaryID = primarySlotToID[primarySlot];
        const b2o_result* pr = &primaryResults[primarySlot];
        const int32_t* batchSlots = batchIDToSlot + (isPrimary ? primaryID : taskIdx * numRaysPerPrimary);
        float color[4] = { 0, 0, 0, 0 };
        for (int i = 0; i < numRaysPerPrimary; i++) {
            int tri = batchResults[batchSlots[i]].id;
            float add[4];
            if (tri == -1) {
                if (isPrimary) memcpy(add, bg, sizeof(add));
                else { add[0] = add[1] = add[2] = add[3] = 1.0f; }
            } else {
                if (isAO) { add[0] = add[1] = add[2] = 0.0f; add[3] = 1.0f; }
                else from_abgr(triShadedColor[tri], add);
            }
            for (int k = 0; k < 4; k++) color[k] += add[k];
        }
        float s = 1.0f / (float)numRaysPerPrimary;
        for (int k = 0; k < 4; k++) color[k] *= s;
        int tri = pr->id;
        if (isAO && tri == -1) memcpy(color, bg, sizeof(color));
        if (isDiffuse) {
            float m[4];
            if (tri == -1) memcpy(m, bg, sizeof(m)); else from_abgr(triMaterialColor[tri], m);
            for (int k = 0; k < 4; k++) color[k] *= m[k];
        }
        pixels[primaryID] = to_abgr(color);
    }
}
