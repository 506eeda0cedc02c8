// bvhcache_key.cpp -- the file name Renderer::getCudaBVH gives a cached CudaBVH (rt/cuda/Renderer.cc:176-186):
//     <dir>/%08x.dat,  key = hashBits(scene.hash(), platform.computeHash(), buildParams.computeHash(), layout)
// restated so that a cache directory can be shared with a reference build: a file the reference wrote is found under the
// same name (and b2rt_bvh_load reads its bytes, host_util.cpp), and a file written here is found by the reference.
//   scene hash      Scene.cc:94-102: hashBits over hashBuffer of the triangle vertex indices, triangle normals, material
//                   colours, shaded colours and vertex positions -- exactly the arrays Scene::Scene builds (Scene.cc:35-84):
//                   normal = normalize(cross(v1 - v0, v2 - v0)), material colour = diffuse.toABGR(), shaded colour =
//                   (diffuse.xyz * (dot(normal, light) * 0.5 + 0.5), 1).toABGR() with light = normalize(1, 2, 3)
//   platform hash   bvh/Platform.hh:69: name ("GPU", Renderer.cc:53), SAH costs 1 / 1, batch sizes 1 / 1, leaf sizes (1, 8)
//   build params    bvh/BVH.hh:82-85: hashBits(floatToBits(splitAlpha))
//   hash functions  framework/base/Hash.hh:181-196 (Jenkins mix), Hash.cc:34-105 (hashBuffer / hashBufferAlign)
// The arrays must be the ones the REFERENCE'S importer produces for the mesh (vertices in order of first use by a face,
// MeshWavefrontIO.cc): tests/golden holds them for the two fixture scenes, with the keys the reference's own code computed.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "../../../include/b2rt.h"

namespace {

typedef uint32_t U32;
const U32 kMagic = 0x9e3779b9u;

inline void mix(U32& a, U32& b, U32& c)
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
inline U32 hashBits3(U32 a, U32 b = kMagic, U32 c = 0) { c += kMagic; mix(a, b, c); return c; }
inline U32 hashBits6(U32 a, U32 b, U32 c, U32 d, U32 e = 0, U32 f = 0) { c += kMagic; mix(a, b, c); a += d; b += e; c += f; mix(a, b, c); return c; }

// every buffer hashed here is 4-byte aligned with a size that is a multiple of 4: Hash.cc's aligned path
U32 hashWords(const void* ptr, size_t bytes)
{
    const U32* src = static_cast<const U32*>(ptr);
    U32 a = kMagic, b = kMagic, c = kMagic;
    int size = (int)bytes;
    while (size >= 12) { a += src[0]; b += src[1]; c += src[2]; mix(a, b, c); src += 3; size -= 12; }
    switch (size) { case 8: b += src[1]; /* fall through */ case 4: a += src[0]; default: break; }
    c += (U32)size;
    mix(a, b, c);
    return c;
}
// hashBuffer of a byte string whose length need not be a multiple of 4 (the platform's name)
U32 hashBytes(const unsigned char* src, int size)
{
    U32 a = kMagic, b = kMagic, c = kMagic;
    while (size >= 12) {
        a += src[0] + (src[1] << 8) + (src[2] << 16) + ((U32)src[3] << 24);
        b += src[4] + (src[5] << 8) + (src[6] << 16) + ((U32)src[7] << 24);
        c += src[8] + (src[9] << 8) + (src[10] << 16) + ((U32)src[11] << 24);
        mix(a, b, c); src += 12; size -= 12;
    }
    switch (size) {
    case 11: c += (U32)src[10] << 16; /* fall through */
    case 10: c += (U32)src[9] << 8;   /* fall through */
    case 9:  c += src[8];             /* fall through */
    case 8:  b += (U32)src[7] << 24;  /* fall through */
    case 7:  b += (U32)src[6] << 16;  /* fall through */
    case 6:  b += (U32)src[5] << 8;   /* fall through */
    case 5:  b += src[4];             /* fall through */
    case 4:  a += (U32)src[3] << 24;  /* fall through */
    case 3:  a += (U32)src[2] << 16;  /* fall through */
    case 2:  a += (U32)src[1] << 8;   /* fall through */
    case 1:  a += src[0];             /* fall through */
    default: break;
    }
    c += (U32)size;
    mix(a, b, c);
    return c;
}
inline U32 floatBits(float f) { U32 u; memcpy(&u, &f, 4); return u; }

// Vec4f::toABGR, one channel (framework/base/Math.cc:46-53)
inline U32 chan(float x)
{
    const float cl = x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x);
    return (U32)((((unsigned long long)(cl * 72057594037927936.0f) * 255ull) >> 55) + 1ull) >> 1;
}
inline U32 toABGR(float x, float y, float z, float w) { return chan(x) | (chan(y) << 8) | (chan(z) << 16) | (chan(w) << 24); }

}  // namespace

extern "C" {

int b2rt_bvhcache_key(const float* verts, int numVerts, const int32_t* tris, int numTris, const float* triDiffuseRGBA,
                      float splitAlpha, int minLeaf, int maxLeaf, int layout, uint32_t* key)
{
    if (!verts || !tris || !key || numVerts < 0 || numTris < 0) return B2RT_ERR_ARG;
    std::vector<float> normals((size_t)numTris * 3);
    int rc = b2rt_tri_normals(verts, tris, numTris, normals.data());      // Scene.cc:72, bit-exact (tests/test_oracle_golden.py)
    if (rc) return rc;
    // light = Vec3f(1, 2, 3).normalized(): v * (1 * rcp(length)), Math.hh:142
    const float len = sqrtf(1.0f * 1.0f + 2.0f * 2.0f + 3.0f * 3.0f);
    const float inv = 1.0f * (1.0f / len);
    const float light[3] = { 1.0f * inv, 2.0f * inv, 3.0f * inv };
    std::vector<U32> material((size_t)numTris), shaded((size_t)numTris);
    for (int i = 0; i < numTris; i++) {
        // MeshBase::Material's default diffuse (3d/Mesh.hh:92) unless the caller has per-triangle colours
        const float dflt[4] = { 0.75f, 0.75f, 0.75f, 1.0f };
        const float* d = triDiffuseRGBA ? triDiffuseRGBA + (size_t)i * 4 : dflt;
        const float* n = &normals[(size_t)i * 3];
        material[i] = toABGR(d[0], d[1], d[2], d[3]);
        const float s = (n[0] * light[0] + n[1] * light[1] + n[2] * light[2]) * 0.5f + 0.5f;
        shaded[i] = toABGR(d[0] * s, d[1] * s, d[2] * s, 1.0f);
    }
    const U32 sceneHash = hashBits6(hashWords(tris, (size_t)numTris * 12), hashWords(normals.data(), (size_t)numTris * 12),
                                    hashWords(material.data(), (size_t)numTris * 4), hashWords(shaded.data(), (size_t)numTris * 4),
                                    hashWords(verts, (size_t)numVerts * 12));
    // Platform("GPU") with setLeafPreferences(minLeaf, maxLeaf): hashBits(hash<String>(name), bits(1.f), bits(1.f), hashBits(1, 1, min, max))
    const U32 platformHash = hashBits6(hashBytes(reinterpret_cast<const unsigned char*>("GPU"), 3), floatBits(1.0f), floatBits(1.0f),
                                       hashBits6(1u, 1u, (U32)minLeaf, (U32)maxLeaf));
    const U32 paramsHash = hashBits3(floatBits(splitAlpha));
    *key = hashBits6(sceneHash, platformHash, paramsHash, (U32)layout);
    return B2RT_OK;
}

int b2rt_bvhcache_path(const char* dir, uint32_t key, char* out, size_t cap)
{
    if (!dir || !out) return B2RT_ERR_ARG;
    const int n = snprintf(out, cap, "%s/%08x.dat", dir, key);            // Renderer.cc:180
    return (n < 0 || (size_t)n >= cap) ? B2RT_ERR_ARG : B2RT_OK;
}

}  // extern "C"
