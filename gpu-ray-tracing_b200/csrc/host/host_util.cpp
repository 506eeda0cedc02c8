// host_util.cpp -- host-side pieces of the ray pipeline that need no GPU.
//
//  * b2rt_pixel_table: the index -> pixel order of primary rays.  Same permutation as
//    PixelTable::recalculate (ray/PixelTable.cc:94-156): 8x8 pixel blocks visited along a Z-order
//    curve over the block grid, Z-order again inside a block, then the bottom stripe (column by
//    column) and the right stripe with the corner (row by row).  This order is what makes
//    consecutive primary rays coherent, so it is part of the hot path's contract.
//  * b2rt_camera_matrices: origin and the normalised-screen -> world matrix of a camera given as
//    position / forward / up / fov / near / far, i.e. what Renderer::beginFrame passes to
//    RayGen::primary (cuda/Renderer.cc:126-129; framework/3d/CameraControls.cc:263-297;
//    framework/base/Math.cc:64-91).  Column-major 4x4, like FW::Mat4f.
// Built with -ffp-contract=off.
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "../../../include/b2rt.h"

namespace {

// de-interleave the even bits of v (x of a Z-order index); y = evenBits(v >> 1)
inline uint32_t evenBits(uint64_t v)
{
    v &= 0x5555555555555555ull;
    v = (v | (v >> 1)) & 0x3333333333333333ull;
    v = (v | (v >> 2)) & 0x0f0f0f0f0f0f0f0full;
    v = (v | (v >> 4)) & 0x00ff00ff00ff00ffull;
    v = (v | (v >> 8)) & 0x0000ffff0000ffffull;
    v = (v | (v >> 16)) & 0x00000000ffffffffull;
    return (uint32_t)v;
}

struct Mat4 {
    float m[16];    // m[r + 4c]
    float& at(int r, int c) { return m[r + 4 * c]; }
    float at(int r, int c) const { return m[r + 4 * c]; }
    static Mat4 identity() { Mat4 r; memset(r.m, 0, sizeof(r.m)); for (int i = 0; i < 4; i++) r.at(i, i) = 1.0f; return r; }
};

Mat4 mul(const Mat4& a, const Mat4& b)
{
    Mat4 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float s = 0.0f;
            for (int k = 0; k < 4; k++) s += a.at(i, k) * b.at(k, j);
            r.at(i, j) = s;
        }
    return r;
}

inline float rcpOrZero(float a) { return (a != 0.0f) ? 1.0f / a : 0.0f; }

inline float det3(const float s[3][3])
{
    return s[0][0] * s[1][1] * s[2][2] - s[0][0] * s[1][2] * s[2][1] +
           s[1][0] * s[2][1] * s[0][2] - s[1][0] * s[2][2] * s[0][1] +
           s[2][0] * s[0][1] * s[1][2] - s[2][0] * s[0][2] * s[1][1];
}

// cofactor inverse in the operation order of MatrixBase::inverted (Math.hh:962-983)
Mat4 inverse(const Mat4& m)
{
    Mat4 r;
    float d = 0.0f, si = 1.0f;
    for (int i = 0; i < 4; i++) {
        float sj = si;
        for (int j = 0; j < 4; j++) {
            float sub[3][3];
            for (int k = 0; k < 3; k++)
                for (int l = 0; l < 3; l++)
                    sub[k][l] = m.at((k < j) ? k : k + 1, (l < i) ? l : l + 1);
            const float dd = det3(sub) * sj;
            r.at(i, j) = dd;
            d += dd * m.at(j, i);
            sj = -sj;
        }
        si = -si;
    }
    const float rd = rcpOrZero(d);
    for (int k = 0; k < 16; k++) r.m[k] = r.m[k] * rd * 4.0f;
    return r;
}

struct V3 { float x, y, z; };
inline float dot(V3 a, V3 b) { float r = 0.0f; r += a.x * b.x; r += a.y * b.y; r += a.z * b.z; return r; }
inline V3 cross(V3 a, V3 b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
inline V3 normalized(V3 a) { const float s = 1.0f * rcpOrZero(sqrtf(dot(a, a))); return { a.x * s, a.y * s, a.z * s }; }

}  // namespace

extern "C" {

int b2rt_pixel_table(int w, int h, int32_t* indexToPixel, int32_t* pixelToIndex)
{
    if (w < 0 || h < 0 || (long long)w * h > 0x7fffffffLL || (!indexToPixel && w * h)) return B2RT_ERR_ARG;
    const int bw = w & ~7, bh = h & ~7;             // part of the image covered by whole 8x8 blocks
    const int blocksX = bw >> 3, blocksY = bh >> 3;
    int side = 1;                                   // power-of-two side of the Z-order curve over blocks
    while (side < blocksX || side < blocksY) side <<= 1;
    int idx = 0;
    auto put = [&](int pos) { if (pixelToIndex) pixelToIndex[pos] = idx; indexToPixel[idx++] = pos; };
    if (blocksX > 0 && blocksY > 0)
        for (uint64_t z = 0; z < (uint64_t)side * side; z++) {
            const int tx = (int)evenBits(z), ty = (int)evenBits(z >> 1);
            if (tx >= blocksX || ty >= blocksY) continue;
            for (uint32_t in = 0; in < 64; in++) {
                const int ix = (int)evenBits(in), iy = (int)evenBits(in >> 1);
                put((ty * 8 + iy) * w + tx * 8 + ix);
            }
        }
    for (int px = 0; px < bw; px++)
        for (int py = bh; py < h; py++) put(px + py * w);
    for (int py = 0; py < h; py++)
        for (int px = bw; px < w; px++) put(px + py * w);
    return B2RT_OK;
}

int b2rt_camera_matrices(const float pos[3], const float forward[3], const float up[3], float fov, float znear, float zfar,
                         int w, int h, float origin[3], float nscreenToWorld[16])
{
    if (!pos || !forward || !up || !origin || !nscreenToWorld || w <= 0 || h <= 0) return B2RT_ERR_ARG;
    const V3 p = { pos[0], pos[1], pos[2] };
    V3 z = normalized({ forward[0], forward[1], forward[2] });
    z = { -z.x, -z.y, -z.z };
    const V3 x = normalized(cross({ up[0], up[1], up[2] }, z));
    const V3 y = normalized(cross(z, x));

    Mat4 worldToCamera = Mat4::identity();
    const V3 axes[3] = { x, y, z };
    for (int r = 0; r < 3; r++) {
        worldToCamera.at(r, 0) = axes[r].x; worldToCamera.at(r, 1) = axes[r].y; worldToCamera.at(r, 2) = axes[r].z;
        worldToCamera.at(r, 3) = -dot(axes[r], p);
    }
    Mat4 persp; memset(persp.m, 0, sizeof(persp.m));
    const float f = rcpOrZero(tanf(fov * 3.14159265358979323846f / 360.0f));
    const float d = rcpOrZero(znear - zfar);
    persp.at(0, 0) = f; persp.at(1, 1) = f;
    persp.at(2, 2) = (znear + zfar) * d; persp.at(2, 3) = 2.0f * znear * zfar * d;
    persp.at(3, 2) = -1.0f;
    const Mat4 worldToClip = mul(persp, worldToCamera);

    // fitToView(pos = -1, size = 2, viewSize = (w, h))
    Mat4 s1 = Mat4::identity(), s2 = Mat4::identity(), tr = Mat4::identity();
    s1.at(0, 0) = 2.0f / (float)w; s1.at(1, 1) = 2.0f / (float)h;
    const float sx = (float)w / 2.0f, sy = (float)h / 2.0f;
    const float mn = (sx < sy) ? sx : sy;
    s2.at(0, 0) = mn; s2.at(1, 1) = mn;
    tr.at(0, 3) = 1.0f - 2.0f * 0.5f; tr.at(1, 3) = 1.0f - 2.0f * 0.5f;
    const Mat4 fit = mul(mul(s1, s2), tr);
    const Mat4 inv = inverse(mul(fit, worldToClip));
    memcpy(nscreenToWorld, inv.m, sizeof(inv.m));
    origin[0] = pos[0]; origin[1] = pos[1]; origin[2] = pos[2];
    return B2RT_OK;
}

}  // extern "C"
