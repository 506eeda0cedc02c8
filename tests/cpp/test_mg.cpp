// test_mg.cpp -- a C++ host driving several GPUs through include/b2rt_mg.h: scene once, rays generated per device through
// the per-device contexts (each device its own range of the frame's primary rays), sharded trace, gathered hit records,
// summed hit count -- compared with the single-device path (b2rt_trace on device 0 over the whole frame).
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "b2rt_mg.h"
#include "b2rt_scenes.h"

#define CHECK(cond, ...) do { if (!(cond)) { printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); return 1; } } while (0)
#define OK(call) do { int rc_ = (call); CHECK(rc_ == 0, "%s -> %d (%s)", #call, rc_, mg ? b2rt_mg_last_error(mg) : b2rt_mg_last_error(nullptr)); } while (0)

struct Result { int id; float t; int padA, padB; };

int main()
{
    b2rt_mg* mg = nullptr;
    int ndev = 0;
    cudaGetDeviceCount(&ndev);
    if (ndev > 2) ndev = 2;
    CHECK(ndev >= 1, "no CUDA device");
    OK(b2rt_mg_create(ndev, nullptr, &mg));
    CHECK(b2rt_mg_num_devices(mg) == ndev, "device count");

    // scene + SBVH on the host, replicated by ONE broadcast call
    b2rt_scene* scene = nullptr;
    CHECK(b2rt_scene_generate(1 /* interior */, 40000, 23, &scene) == 0, "scene");
    const float* verts; const int32_t* tris; int nv, nt;
    b2rt_scene_arrays(scene, &verts, &nv, &tris, &nt);
    float cam[12];
    b2rt_scene_camera(scene, cam);
    b2rt_host_bvh* hbvh = nullptr;
    CHECK(b2rt_bvh_build(verts, nv, tris, nt, nullptr, &hbvh) == 0, "bvh build");
    const void *nodes, *woop; const int32_t* triIndex; size_t nb, wb, tb;
    b2rt_bvh_arrays(hbvh, &nodes, &nb, &woop, &wb, &triIndex, &tb);
    OK(b2rt_mg_broadcast_bvh(mg, nodes, nb, woop, wb, triIndex, tb, B2RT_LAYOUT_COMPACT2, 0));

    // the frame's primary rays, generated on device 0, then handed out in two unequal shards
    const int W = 640, H = 480, N = W * H;
    float origin[3], m[16];
    CHECK(b2rt_camera_matrices(cam, cam + 3, cam + 6, cam[9], cam[10], cam[11], W, H, origin, m) == 0, "camera");
    std::vector<int32_t> i2p(N);
    b2rt_pixel_table(W, H, i2p.data(), nullptr);
    cudaSetDevice(0);
    int32_t *dI2p, *dA, *dB; void *dRays, *dWant;
    cudaMalloc(&dI2p, N * 4); cudaMalloc(&dA, N * 4); cudaMalloc(&dB, N * 4); cudaMalloc(&dRays, (size_t)N * 32); cudaMalloc(&dWant, (size_t)N * 16);
    cudaMemcpy(dI2p, i2p.data(), N * 4, cudaMemcpyHostToDevice);
    b2rt_ctx* c0 = b2rt_mg_context(mg, 0);
    CHECK(b2rt_raygen_primary(c0, origin, m, W, H, cam[11], dI2p, dRays, dA, dB, nullptr) == 0, "raygen: %s", b2rt_last_error(c0));
    float ms = 0;
    CHECK(b2rt_trace(c0, dRays, dWant, N, 0, nullptr, &ms) == 0, "single-device trace: %s", b2rt_last_error(c0));
    std::vector<Result> want(N);
    cudaMemcpy(want.data(), dWant, (size_t)N * 16, cudaMemcpyDeviceToHost);
    std::vector<char> hostRays((size_t)N * 32);
    cudaMemcpy(hostRays.data(), dRays, (size_t)N * 32, cudaMemcpyDeviceToHost);

    int64_t count[2] = { ndev == 2 ? (int64_t)N * 2 / 3 + 5 : N, 0 };
    if (ndev == 2) count[1] = N - count[0];
    const void* rays[2]; void* results[2]; void* gathered[2];
    for (int r = 0; r < ndev; r++) {
        cudaSetDevice(r);
        void *pr, *ps, *pg;
        cudaMalloc(&pr, (size_t)count[r] * 32 + 32); cudaMalloc(&ps, (size_t)count[r] * 16 + 16); cudaMalloc(&pg, (size_t)N * 16);
        cudaMemcpy(pr, hostRays.data() + (r ? (size_t)count[0] * 32 : 0), (size_t)count[r] * 32, cudaMemcpyHostToDevice);
        cudaMemset(ps, 0, (size_t)count[r] * 16 + 16);
        cudaMemset(pg, 0xff, (size_t)N * 16);
        rays[r] = pr; results[r] = ps; gathered[r] = pg;
    }
    b2rt_mg_stats st;
    OK(b2rt_mg_trace(mg, rays, results, count, 0, gathered, &st));
    long long wantHits = 0;
    for (int i = 0; i < N; i++) wantHits += want[i].id >= 0;
    CHECK(st.rays == N && st.hits == wantHits, "stats: rays %lld hits %lld (want %d, %lld)", (long long)st.rays, (long long)st.hits, N, wantHits);
    CHECK(st.ms_trace_max > 0.0f && st.ms_with_gather_max >= st.ms_trace_max, "times %g %g", st.ms_trace_max, st.ms_with_gather_max);
    std::vector<Result> got(N);
    for (int r = 0; r < ndev; r++) {
        cudaSetDevice(r);
        cudaMemcpy(got.data(), gathered[r], (size_t)N * 16, cudaMemcpyDeviceToHost);
        int same = 0;
        for (int i = 0; i < N; i++) {
            if (got[i].id == want[i].id) { same++; CHECK(got[i].t == want[i].t, "device %d ray %d: t %g vs %g", r, i, got[i].t, want[i].t); }
        }
        CHECK(same >= (long long)N * 9999 / 10000, "device %d: %d of %d ids agree", r, same, N);
    }
    // host buffers
    std::vector<Result> out(N);
    b2rt_mg_stats st2;
    OK(b2rt_mg_trace_host(mg, hostRays.data(), out.data(), N, 0, &st2));
    CHECK(st2.hits == wantHits, "host path hits %lld", (long long)st2.hits);
    int same = 0;
    for (int i = 0; i < N; i++) same += out[i].id == want[i].id;
    CHECK(same >= (long long)N * 9999 / 10000, "host path: %d of %d ids agree", same, N);
    // errors: a NULL shard with a count is refused, the group stays usable
    const void* badRays[2] = { nullptr, nullptr };
    CHECK(b2rt_mg_trace(mg, badRays, results, count, 0, nullptr, nullptr) != 0, "NULL shard accepted");
    CHECK(strstr(b2rt_mg_last_error(mg), "bad shard") != nullptr, "message: %s", b2rt_mg_last_error(mg));
    OK(b2rt_mg_trace(mg, rays, results, count, 0, nullptr, &st));
    printf("test_mg ok: %d device(s), %d rays, %lld hits, trace %.3f ms (max over devices), with gather %.3f ms\n", ndev, N, wantHits,
           st.ms_trace_max, st2.ms_trace_max);
    b2rt_mg_destroy(mg);
    b2rt_bvh_free(hbvh);
    b2rt_scene_free(scene);
    return 0;
}
