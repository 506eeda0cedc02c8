// test_fw.cpp -- the reference's call sequence (App.cc:137-204, Renderer.cc:112-300) written against
// include/b2rt_fw.hpp, checked against the oracle (test infrastructure) on the same rays and BVH.
//   build: tests/test_cpp_shim.py;  run: ./test_fw  (needs a B200)
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "b2rt_fw.hpp"
#include "b2rt_scenes.h"
#include "b2rt_oracle.h"

namespace FW = b2rt_fw;

#define CHECK(cond, ...) do { if (!(cond)) { printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); return 1; } } while (0)

int main()
{
    // scene + SBVH (host)
    b2rt_scene* scene = nullptr;
    FW::check(b2rt_scene_generate(1 /* interior */, 30000, 11, &scene));
    const float* verts; const int32_t* tris; int nv, nt;
    FW::check(b2rt_scene_arrays(scene, &verts, &nv, &tris, &nt));
    float cam[12];
    FW::check(b2rt_scene_camera(scene, cam));
    b2rt_host_bvh* hbvh = nullptr;
    FW::check(b2rt_bvh_build(verts, nv, tris, nt, nullptr, &hbvh));
    const void *nodes, *woop; const int32_t* triIndex; size_t nb, wb, tb;
    FW::check(b2rt_bvh_arrays(hbvh, &nodes, &nb, &woop, &wb, &triIndex, &tb));
    std::vector<float> normals((size_t)nt * 3);
    FW::check(b2rt_tri_normals(verts, tris, nt, normals.data()));

    // Renderer::beginFrame
    FW::CudaTracer tracer;
    FW::RayBuffer empty;
    CHECK(tracer.traceBatch(empty) == 0.0f, "empty batch must return 0 ms");
    FW::RayBuffer four(4);
    bool threw = false;
    try { tracer.traceBatch(four); } catch (const std::runtime_error& e) { threw = strcmp(e.what(), "CudaTracer: No BVH!") == 0; }
    CHECK(threw, "missing BVH must fail with the reference's message");
    FW::CudaBVH bvh(hbvh);
    tracer.setBVH(&bvh);
    const int W = 256, H = 192;
    float origin[3], m[16];
    FW::check(b2rt_camera_matrices(cam, cam + 3, cam + 6, cam[9], cam[10], cam[11], W, H, origin, m));
    FW::RayGen raygen(1 << 21);
    raygen.setTriNormals(normals.data(), nt);
    FW::RayBuffer primary;
    raygen.primary(primary, FW::Vec3f{ origin[0], origin[1], origin[2] }, m, W, H, cam[11]);
    float ms = tracer.traceBatch(primary);
    CHECK(ms > 0.0f, "kernel time");

    // parity of the primary batch against the oracle's scalar restatement of the GPU algorithm
    std::vector<FW::Ray> rays = primary.getRays();
    std::vector<FW::RayResult> got = primary.getResults();
    std::vector<b2o_result> ref(rays.size());
    b2o_trace_woop((const float*)nodes, (const float*)woop, triIndex, (const b2o_ray*)rays.data(), ref.data(), (int)rays.size(), 0, nullptr);
    size_t same = 0, hits = 0;
    double maxRel = 0.0;
    for (size_t i = 0; i < rays.size(); i++) {
        if (got[i].id == ref[i].id) {
            same++;
            if (ref[i].id >= 0) { double r = fabs(got[i].t - ref[i].t) / fabs(ref[i].t); if (r > maxRel) maxRel = r; }
            else CHECK(got[i].t == rays[i].tmax, "miss must return t = tmax");
        }
        if (got[i].id >= 0) hits++;
    }
    CHECK(same >= (size_t)(0.9999 * rays.size()), "primary id agreement %zu / %zu", same, rays.size());
    CHECK(maxRel <= 1e-5, "primary t error %g", maxRel);
    CHECK((size_t)FW::countHits(primary) == hits, "countHits");

    // Renderer::nextBatch, Diffuse: closest hit, then AO: sorted, any hit
    FW::RayBuffer secondary;
    bool newBatch = true;
    CHECK(raygen.ao(secondary, primary, 2, cam[11], newBatch), "ao batch");
    secondary.setNeedClosestHit(true);
    tracer.traceBatch(secondary);
    rays = secondary.getRays(); got = secondary.getResults(); ref.resize(rays.size());
    b2o_trace_woop((const float*)nodes, (const float*)woop, triIndex, (const b2o_ray*)rays.data(), ref.data(), (int)rays.size(), 0, nullptr);
    same = 0;
    for (size_t i = 0; i < rays.size(); i++) same += got[i].id == ref[i].id;
    CHECK(same >= (size_t)(0.9999 * rays.size()), "diffuse id agreement %zu / %zu", same, rays.size());

    newBatch = true;
    CHECK(raygen.ao(secondary, primary, 4, 1.0f, newBatch), "ao batch");
    std::vector<FW::Ray> before = secondary.getRays();
    secondary.mortonSort();
    rays = secondary.getRays();
    std::vector<FW::S32> s2i = secondary.getSlotToID(), i2s = secondary.getIDToSlot();
    for (size_t i = 0; i < rays.size(); i++) {
        CHECK(memcmp(&rays[i], &before[s2i[i]], sizeof(FW::Ray)) == 0, "sorted slot %zu does not hold ray id %d", i, s2i[i]);
        CHECK(i2s[s2i[i]] == (FW::S32)i, "id maps are not inverse");
    }
    tracer.traceBatch(secondary);
    got = secondary.getResults(); ref.resize(rays.size());
    b2o_trace_woop((const float*)nodes, (const float*)woop, triIndex, (const b2o_ray*)rays.data(), ref.data(), (int)rays.size(), 1, nullptr);
    same = 0;
    for (size_t i = 0; i < rays.size(); i++) same += (got[i].id >= 0) == (ref[i].id >= 0);
    CHECK(same >= (size_t)(0.9999 * rays.size()), "any-hit agreement %zu / %zu", same, rays.size());

    b2rt_bvh_free(hbvh);
    b2rt_scene_free(scene);
    printf("test_fw ok: %d primary rays in %.3f ms, t error %.2e\n", W * H, ms, maxRel);
    return 0;
}
