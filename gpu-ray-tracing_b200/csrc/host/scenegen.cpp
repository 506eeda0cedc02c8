// scenegen.cpp -- seeded procedural stand-ins for the scenes the reference is quoted on.
//
// The README scenes (bunny 144 500 tris, conference 350 949, dragon 910 348, hairball 6 469 561;
// README.md:48-58) are not in the reference tree and cannot be fetched (no network), so the
// benchmark configs use generated meshes of the same triangle counts and the same character:
//   kind 0 "blob"     closed genus-0 surface with low-frequency radial displacement (bunny-scale)
//   kind 1 "interior" axis-aligned room full of tables and chairs: large flat walls next to many
//                     small parts -- the case spatial splits exist for (conference-scale)
//   kind 2 "knot"     long thin closed tube tied in a torus knot, finely tessellated (dragon-scale)
//   kind 3 "strands"  ~50 K thin curved 3-sided tubes growing out of a ball (hairball-scale)
// Generator = SplitMix64; triangle counts hit the target exactly (the last grid row is trimmed
// or a few triangles are appended), vertices are shared inside a grid.  Cameras follow
// CameraControls::initForMesh (framework/3d/CameraControls.cc:330-350) except for the interior.
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include <algorithm>
#include <vector>

#include "../../../include/b2rt_scenes.h"

#define B2RT_OK 0
#define B2RT_ERR_ARG (-1)

struct b2rt_scene {
    std::vector<float> verts;
    std::vector<int32_t> tris;
    float cam[9];       // position, forward, up
    float fov, znear, zfar;
};

namespace {

struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed) {}
    uint64_t next() { uint64_t z = (s += 0x9e3779b97f4a7c15ull); z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull; z = (z ^ (z >> 27)) * 0x94d049bb133111ebull; return z ^ (z >> 31); }
    float uni() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }          // [0,1)
    float range(float a, float b) { return a + (b - a) * uni(); }
};

struct Mesh {
    std::vector<float>& v; std::vector<int32_t>& t;
    int addVertex(float x, float y, float z) { v.push_back(x); v.push_back(y); v.push_back(z); return (int)(v.size() / 3) - 1; }
    void addTri(int a, int b, int c) { t.push_back(a); t.push_back(b); t.push_back(c); }
    size_t numTris() const { return t.size() / 3; }
};

const float kPi = 3.14159265358979323846f;

// low-frequency smooth field on R^3: a fixed sum of seeded sinusoids
struct Field {
    float a[12][4];
    explicit Field(Rng& r, float freq)
    {
        for (int i = 0; i < 12; i++) {
            const float f = freq * (1.0f + (float)(i / 4));
            a[i][0] = r.range(-f, f); a[i][1] = r.range(-f, f); a[i][2] = r.range(-f, f); a[i][3] = r.range(0.0f, 2.0f * kPi);
        }
    }
    float eval(float x, float y, float z) const
    {
        float s = 0.0f;
        for (int i = 0; i < 12; i++) s += sinf(a[i][0] * x + a[i][1] * y + a[i][2] * z + a[i][3]) / (float)(1 + i / 4);
        return s * (1.0f / 6.0f);
    }
};

// pick grid dimensions nu x nv with 2 nu nv >= target and nu/nv ~ aspect
void gridFor(long long target, float aspect, int& nu, int& nv)
{
    const double quads = (double)target / 2.0;
    nv = std::max(3, (int)floor(sqrt(quads / aspect)));
    nu = std::max(3, (int)((quads + nv - 1) / nv));
    while ((long long)2 * nu * nv < target) nu++;
}

// closed (u wraps, v wraps or pole-capped) parametric grid; emits exactly `target` triangles by
// dropping triangles from the end of the last row
template <class F>
void emitGrid(Mesh& m, int nu, int nv, bool wrapU, bool wrapV, long long target, F pos)
{
    const int cu = wrapU ? nu : nu + 1, cv = wrapV ? nv : nv + 1;
    const int base = (int)(m.v.size() / 3);
    for (int i = 0; i < cu; i++)
        for (int j = 0; j < cv; j++) {
            float p[3];
            pos((float)i / (float)nu, (float)j / (float)nv, p);
            m.addVertex(p[0], p[1], p[2]);
        }
    auto id = [&](int i, int j) { return base + (i % cu) * cv + (j % cv); };
    long long emitted = 0;
    for (int i = 0; i < nu && emitted < target; i++)
        for (int j = 0; j < nv && emitted < target; j++) {
            const int a = id(i, j), b = id(i + 1, j), c = id(i + 1, j + 1), d = id(i, j + 1);
            m.addTri(a, b, c); emitted++;
            if (emitted < target) { m.addTri(a, c, d); emitted++; }
        }
}

void addBox(Mesh& m, const float lo[3], const float hi[3])
{
    int id[8];
    for (int k = 0; k < 8; k++) id[k] = m.addVertex((k & 1) ? hi[0] : lo[0], (k & 2) ? hi[1] : lo[1], (k & 4) ? hi[2] : lo[2]);
    const int f[6][4] = { {0,2,3,1}, {4,5,7,6}, {0,1,5,4}, {2,6,7,3}, {0,4,6,2}, {1,3,7,5} };
    for (int k = 0; k < 6; k++) { m.addTri(id[f[k][0]], id[f[k][1]], id[f[k][2]]); m.addTri(id[f[k][0]], id[f[k][2]], id[f[k][3]]); }
}

// open cylinder along +y with `sides` sides and `rings` rings: 2*sides*rings triangles
void addCylinder(Mesh& m, float cx, float y0, float cz, float radius, float height, int sides, int rings)
{
    const int base = (int)(m.v.size() / 3);
    for (int r = 0; r <= rings; r++)
        for (int s = 0; s < sides; s++) {
            const float a = 2.0f * kPi * (float)s / (float)sides;
            m.addVertex(cx + radius * cosf(a), y0 + height * (float)r / (float)rings, cz + radius * sinf(a));
        }
    for (int r = 0; r < rings; r++)
        for (int s = 0; s < sides; s++) {
            const int a = base + r * sides + s, b = base + r * sides + (s + 1) % sides, c = a + sides, d = b + sides;
            m.addTri(a, b, d); m.addTri(a, d, c);
        }
}

void addQuadGrid(Mesh& m, const float o[3], const float eu[3], const float ev[3], int nu, int nv)
{
    const int base = (int)(m.v.size() / 3);
    for (int i = 0; i <= nu; i++)
        for (int j = 0; j <= nv; j++) {
            const float fu = (float)i / (float)nu, fv = (float)j / (float)nv;
            m.addVertex(o[0] + eu[0] * fu + ev[0] * fv, o[1] + eu[1] * fu + ev[1] * fv, o[2] + eu[2] * fu + ev[2] * fv);
        }
    for (int i = 0; i < nu; i++)
        for (int j = 0; j < nv; j++) {
            const int a = base + i * (nv + 1) + j, b = a + (nv + 1), c = b + 1, d = a + 1;
            m.addTri(a, b, c); m.addTri(a, c, d);
        }
}

void setCameraForBounds(b2rt_scene* s)
{
    float lo[3] = { 3e38f, 3e38f, 3e38f }, hi[3] = { -3e38f, -3e38f, -3e38f };
    for (size_t i = 0; i < s->verts.size(); i += 3)
        for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], s->verts[i + k]); hi[k] = std::max(hi[k], s->verts[i + k]); }
    const float size = std::max(hi[0] - lo[0], std::max(hi[1] - lo[1], hi[2] - lo[2]));
    s->cam[0] = (lo[0] + hi[0]) * 0.5f; s->cam[1] = (lo[1] + hi[1]) * 0.5f; s->cam[2] = (lo[2] + hi[2]) * 0.5f + size * 0.75f;
    s->cam[3] = 0.0f; s->cam[4] = 0.0f; s->cam[5] = -1.0f;
    s->cam[6] = 0.0f; s->cam[7] = 1.0f; s->cam[8] = 0.0f;
    s->fov = 70.0f; s->znear = size * 0.0005f; s->zfar = size * 1.5f;
}

void genBlob(b2rt_scene* s, long long target, uint64_t seed)
{
    Rng rng(seed);
    Field f(rng, 2.5f);
    Mesh m{ s->verts, s->tris };
    int nu, nv;
    gridFor(target, 2.0f, nu, nv);
    // latitude rows are pole-capped (v in (0,1)), longitude wraps
    emitGrid(m, nu, nv, true, false, target, [&](float u, float v, float* p) {
        const float th = kPi * (0.002f + 0.996f * v), ph = 2.0f * kPi * u;
        const float dx = sinf(th) * cosf(ph), dy = cosf(th), dz = sinf(th) * sinf(ph);
        const float r = 1.0f + 0.25f * f.eval(dx, dy, dz);
        p[0] = r * dx; p[1] = r * dy; p[2] = r * dz;
    });
    setCameraForBounds(s);
}

void genKnot(b2rt_scene* s, long long target, uint64_t seed)
{
    Rng rng(seed);
    Field f(rng, 6.0f);
    Mesh m{ s->verts, s->tris };
    int nu, nv;
    gridFor(target, 24.0f, nu, nv);
    auto center = [](float u, float* c) {
        const float t = 2.0f * kPi * u;
        const float r = 2.0f + cosf(3.0f * t);
        c[0] = r * cosf(2.0f * t); c[1] = r * sinf(2.0f * t); c[2] = -1.4f * sinf(3.0f * t);
    };
    emitGrid(m, nu, nv, true, true, target, [&](float u, float v, float* p) {
        float c0[3], c1[3];
        center(u, c0); center(u + 1.0e-3f, c1);
        float tx = c1[0] - c0[0], ty = c1[1] - c0[1], tz = c1[2] - c0[2];
        const float tl = 1.0f / sqrtf(tx * tx + ty * ty + tz * tz);
        tx *= tl; ty *= tl; tz *= tl;
        // frame: n = normalize(z x t), b = t x n
        float nx = -ty, ny = tx, nz = 0.0f;
        const float nl = 1.0f / sqrtf(nx * nx + ny * ny + 1.0e-12f);
        nx *= nl; ny *= nl;
        const float bx = ty * nz - tz * ny, by = tz * nx - tx * nz, bz = tx * ny - ty * nx;
        const float a = 2.0f * kPi * v;
        const float ca = cosf(a), sa = sinf(a);
        const float rad = 0.45f * (1.0f + 0.35f * f.eval(c0[0] + ca, c0[1] + sa, c0[2] + v));
        p[0] = c0[0] + rad * (ca * nx + sa * bx);
        p[1] = c0[1] + rad * (ca * ny + sa * by);
        p[2] = c0[2] + rad * (ca * nz + sa * bz);
    });
    setCameraForBounds(s);
}

void genStrands(b2rt_scene* s, long long target, uint64_t seed)
{
    Rng rng(seed);
    Mesh m{ s->verts, s->tris };
    const int segs = 21, sides = 3, perStrand = segs * sides * 2;
    // a small core ball the strands grow from
    {
        int nu, nv;
        const long long core = std::min<long long>(target / 50, 20000);
        gridFor(core, 2.0f, nu, nv);
        emitGrid(m, nu, nv, true, false, core, [&](float u, float v, float* p) {
            const float th = kPi * (0.002f + 0.996f * v), ph = 2.0f * kPi * u;
            p[0] = 0.3f * sinf(th) * cosf(ph); p[1] = 0.3f * cosf(th); p[2] = 0.3f * sinf(th) * sinf(ph);
        });
    }
    while ((long long)m.numTris() < target) {
        const long long left = target - (long long)m.numTris();
        const int mySegs = (left >= perStrand) ? segs : (int)std::max<long long>(1, left / (sides * 2));
        // random direction on the sphere
        const float z = rng.range(-1.0f, 1.0f), ph = rng.range(0.0f, 2.0f * kPi);
        const float rr = sqrtf(std::max(0.0f, 1.0f - z * z));
        float d[3] = { rr * cosf(ph), z, rr * sinf(ph) };
        float p[3] = { 0.3f * d[0], 0.3f * d[1], 0.3f * d[2] };
        const float step = rng.range(0.03f, 0.05f), radius = rng.range(0.0015f, 0.003f);
        const int base = (int)(m.v.size() / 3);
        for (int k = 0; k <= mySegs; k++) {
            // frame around d
            float ax[3] = { 0.0f, 1.0f, 0.0f };
            if (fabsf(d[1]) > 0.9f) { ax[0] = 1.0f; ax[1] = 0.0f; }
            float n[3] = { d[1] * ax[2] - d[2] * ax[1], d[2] * ax[0] - d[0] * ax[2], d[0] * ax[1] - d[1] * ax[0] };
            const float nl = 1.0f / sqrtf(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
            n[0] *= nl; n[1] *= nl; n[2] *= nl;
            const float b[3] = { d[1] * n[2] - d[2] * n[1], d[2] * n[0] - d[0] * n[2], d[0] * n[1] - d[1] * n[0] };
            const float taper = radius * (1.0f - 0.7f * (float)k / (float)segs);
            for (int sd = 0; sd < sides; sd++) {
                const float a = 2.0f * kPi * (float)sd / (float)sides;
                m.addVertex(p[0] + taper * (cosf(a) * n[0] + sinf(a) * b[0]), p[1] + taper * (cosf(a) * n[1] + sinf(a) * b[1]),
                            p[2] + taper * (cosf(a) * n[2] + sinf(a) * b[2]));
            }
            // random-walk curvature
            d[0] += rng.range(-0.35f, 0.35f); d[1] += rng.range(-0.35f, 0.35f); d[2] += rng.range(-0.35f, 0.35f);
            const float dl = 1.0f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
            d[0] *= dl; d[1] *= dl; d[2] *= dl;
            p[0] += step * d[0]; p[1] += step * d[1]; p[2] += step * d[2];
        }
        for (int k = 0; k < mySegs; k++)
            for (int sd = 0; sd < sides; sd++) {
                if ((long long)m.numTris() >= target) break;
                const int a = base + k * sides + sd, b2 = base + k * sides + (sd + 1) % sides, c = a + sides, d2 = b2 + sides;
                m.addTri(a, b2, d2);
                if ((long long)m.numTris() < target) m.addTri(a, d2, c);
            }
    }
    setCameraForBounds(s);
}

void genInterior(b2rt_scene* s, long long target, uint64_t seed)
{
    Rng rng(seed);
    Mesh m{ s->verts, s->tris };
    const float W = 20.0f, H = 6.0f, D = 12.0f;
    // walls, floor, ceiling: coarse grids (large triangles)
    {
        const float o0[3] = { 0, 0, 0 }, ex[3] = { W, 0, 0 }, ey[3] = { 0, H, 0 }, ez[3] = { 0, 0, D };
        const float o1[3] = { 0, H, 0 }, o2[3] = { W, 0, 0 }, o3[3] = { 0, 0, D };
        addQuadGrid(m, o0, ex, ez, 10, 6);      // floor
        addQuadGrid(m, o1, ez, ex, 6, 10);      // ceiling
        addQuadGrid(m, o0, ey, ex, 3, 10);      // back wall z = 0
        addQuadGrid(m, o3, ex, ey, 10, 3);      // front wall z = D
        addQuadGrid(m, o0, ez, ey, 6, 3);       // left wall
        addQuadGrid(m, o2, ey, ez, 3, 6);       // right wall
    }
    // furniture: a table with 4 chairs is one "set"; sets on a jittered grid until the budget is spent
    const int legSides = 12, legRings = 6;
    auto chair = [&](float cx, float cz, float rot) {
        const float c = cosf(rot), sn = sinf(rot);
        auto leg = [&](float lx, float lz) { addCylinder(m, cx + c * lx - sn * lz, 0.0f, cz + sn * lx + c * lz, 0.02f, 0.45f, legSides, legRings); };
        leg(-0.2f, -0.2f); leg(0.2f, -0.2f); leg(-0.2f, 0.2f); leg(0.2f, 0.2f);
        const float lo[3] = { cx - 0.25f, 0.45f, cz - 0.25f }, hi[3] = { cx + 0.25f, 0.5f, cz + 0.25f };
        addBox(m, lo, hi);
        // back rest: a bent grid
        const float bo[3] = { cx - 0.25f * c + 0.25f * sn, 0.5f, cz - 0.25f * sn - 0.25f * c };
        const float bu[3] = { 0.5f * c, 0.0f, 0.5f * sn }, bv[3] = { 0.02f * -sn, 0.55f, 0.02f * c };
        addQuadGrid(m, bo, bu, bv, 8, 10);
    };
    auto table = [&](float cx, float cz) {
        const float lo[3] = { cx - 0.9f, 0.72f, cz - 0.5f }, hi[3] = { cx + 0.9f, 0.76f, cz + 0.5f };
        addBox(m, lo, hi);
        addCylinder(m, cx - 0.8f, 0.0f, cz - 0.4f, 0.03f, 0.72f, legSides, legRings);
        addCylinder(m, cx + 0.8f, 0.0f, cz - 0.4f, 0.03f, 0.72f, legSides, legRings);
        addCylinder(m, cx - 0.8f, 0.0f, cz + 0.4f, 0.03f, 0.72f, legSides, legRings);
        addCylinder(m, cx + 0.8f, 0.0f, cz + 0.4f, 0.03f, 0.72f, legSides, legRings);
        // a finely tessellated table cloth
        const float o[3] = { cx - 0.7f, 0.765f, cz - 0.35f }, eu[3] = { 1.4f, 0, 0 }, ev[3] = { 0, 0, 0.7f };
        addQuadGrid(m, o, eu, ev, 14, 8);
    };
    int gx = 0, gz = 0;
    const int cols = 8, rows = 5;
    long long guard = 0;
    while ((long long)m.numTris() + 4000 < target && guard++ < 100000) {
        const float cx = 1.6f + (W - 3.2f) * ((float)gx + rng.range(0.2f, 0.8f)) / (float)cols;
        const float cz = 1.4f + (D - 2.8f) * ((float)gz + rng.range(0.2f, 0.8f)) / (float)rows;
        table(cx, cz);
        chair(cx - 0.5f, cz - 0.85f, rng.range(-0.3f, 0.3f));
        chair(cx + 0.5f, cz - 0.85f, rng.range(-0.3f, 0.3f));
        chair(cx - 0.5f, cz + 0.85f, kPi + rng.range(-0.3f, 0.3f));
        chair(cx + 0.5f, cz + 0.85f, kPi + rng.range(-0.3f, 0.3f));
        if (++gx == cols) { gx = 0; if (++gz == rows) gz = 0; }
    }
    // spend what is left on a relief panel hung on the back wall
    {
        const long long left = target - (long long)m.numTris();
        if (left > 0) {
            int nu, nv;
            gridFor(left, 3.0f, nu, nv);
            Field f(rng, 3.0f);
            emitGrid(m, nu, nv, false, false, left, [&](float u, float v, float* p) {
                p[0] = 2.0f + 16.0f * u; p[1] = 1.0f + 4.0f * v; p[2] = 0.05f + 0.04f * f.eval(8.0f * u, 3.0f * v, 0.0f);
            });
        }
    }
    // camera inside the room, looking down the long axis
    s->cam[0] = 1.0f; s->cam[1] = 2.2f; s->cam[2] = D * 0.5f + 0.8f;
    s->cam[3] = 0.96f; s->cam[4] = -0.22f; s->cam[5] = -0.17f;
    s->cam[6] = 0.0f; s->cam[7] = 1.0f; s->cam[8] = 0.0f;
    s->fov = 70.0f; s->znear = 0.01f; s->zfar = 60.0f;
}

}  // namespace

extern "C" {

int b2rt_scene_generate(int kind, int targetTris, uint64_t seed, b2rt_scene** out)
{
    if (!out) return B2RT_ERR_ARG;
    *out = nullptr;
    if (kind < 0 || kind > 3 || targetTris < 64) return B2RT_ERR_ARG;
    b2rt_scene* s = new b2rt_scene();
    switch (kind) {
    case 0: genBlob(s, targetTris, seed); break;
    case 1: genInterior(s, targetTris, seed); break;
    case 2: genKnot(s, targetTris, seed); break;
    default: genStrands(s, targetTris, seed); break;
    }
    *out = s;
    return B2RT_OK;
}

void b2rt_scene_free(b2rt_scene* s) { delete s; }

int b2rt_scene_arrays(const b2rt_scene* s, const float** verts, int* numVerts, const int32_t** tris, int* numTris)
{
    if (!s) return B2RT_ERR_ARG;
    if (verts) *verts = s->verts.data();
    if (numVerts) *numVerts = (int)(s->verts.size() / 3);
    if (tris) *tris = s->tris.data();
    if (numTris) *numTris = (int)(s->tris.size() / 3);
    return B2RT_OK;
}

int b2rt_scene_camera(const b2rt_scene* s, float cam[12])
{
    if (!s || !cam) return B2RT_ERR_ARG;
    for (int k = 0; k < 9; k++) cam[k] = s->cam[k];
    cam[9] = s->fov; cam[10] = s->znear; cam[11] = s->zfar;
    return B2RT_OK;
}

int b2rt_scene_write_obj(const b2rt_scene* s, const char* path)
{
    if (!s || !path) return B2RT_ERR_ARG;
    FILE* f = fopen(path, "w");
    if (!f) return B2RT_ERR_ARG;
    for (size_t i = 0; i < s->verts.size(); i += 3) fprintf(f, "v %.9g %.9g %.9g\n", s->verts[i], s->verts[i + 1], s->verts[i + 2]);
    for (size_t i = 0; i < s->tris.size(); i += 3) fprintf(f, "f %d %d %d\n", s->tris[i] + 1, s->tris[i + 1] + 1, s->tris[i + 2] + 1);
    fclose(f);
    return B2RT_OK;
}

}  // extern "C"
