/*
 * route_a_driver.cc -- proof of drop-in route A: the reference's UNCHANGED host classes on the new kernels.
 *
 * TEST INFRASTRUCTURE (oracle/_ref/route_a).  This file is ours; everything it calls is the reference's own code, compiled from
 * where it lies under /root/reference by oracle/Makefile: CudaTracer::traceBatch (rt/cuda/CudaTracer.cc:119-178 and the
 * owner's dump code behind it, :181-637, with its fetch_tri.cu / fetch_ray.cu / fetch_node.cu), RayGen::primary / ::ao
 * (rt/ray/RayGen.cc), RayBuffer::mortonSort (rt/ray/RayBuffer.cc:256-324), PixelTable, CudaBVH, SplitBVHBuilder, the OBJ importer.
 * What the link replaces is exactly what route A says it replaces: the reference's four kernel files
 * (kepler_dynamic_fetch.cu, RayGenKernels.cu, RayBufferKernels.cu, RendererKernels.cu) are NOT linked; their C-linkage
 * launchers come from libb2rt_refabi.so -> libb2rt.so.  The CUDA runtime and driver are the real ones (no stubs.c); only the
 * GL entry points Buffer.cc names stay link stubs (no graphics stack on this box; they are never reached).
 *
 * Note on the dump code: CudaTracer.cc ships with the triangle dumps switched ON (#define printf_tri_x..w, :50-53), so every
 * traceBatch reads 250 000 triangles (750 000 Woop rows, hard-coded) with a <<<1,1>>> kernel and writes twelve hex files into
 * the working directory.  The scene therefore needs at least 750 000 Woop rows, and the caller runs this in a scratch directory.
 */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

#include "base/Defs.hh"
#include "base/Math.hh"
#include "3d/Mesh.hh"
#include "3d/CameraControls.hh"
#include "io/File.hh"
#include "Scene.hh"
#include "bvh/BVH.hh"
#include "cuda/CudaBVH.hh"
#include "cuda/CudaTracer.hh"
#include "ray/RayBuffer.hh"
#include "ray/RayGen.hh"

using namespace FW;

static void die(const char* msg) { fprintf(stderr, "route_a: %s\n", msg); exit(2); }

static void dump(const std::string& path, const void* dev, size_t count, size_t elem)
{
    std::vector<char> host(count * elem);
    if (cudaMemcpy(host.data(), dev, count * elem, cudaMemcpyDeviceToHost) != cudaSuccess) die("cudaMemcpy D2H failed");
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) die(("cannot write " + path).c_str());
    int n = (int)count;
    fwrite(&n, 4, 1, f);
    fwrite(host.data(), elem, count, f);
    fclose(f);
}

int main(int argc, char** argv)
{
    std::string obj, camSig, out = "route_a";
    int W = 1024, H = 768, samples = 2;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { if (i + 1 >= argc) die("missing argument"); return argv[++i]; };
        if (a == "--obj") obj = next();
        else if (a == "--camera-sig") camSig = next();
        else if (a == "--size") { W = atoi(next()); H = atoi(next()); }
        else if (a == "--samples") samples = atoi(next());
        else if (a == "--out") out = next();
        else die("usage: route_a --obj F --camera-sig S [--size W H] [--samples S] [--out PREFIX]");
    }
    if (obj.empty() || camSig.empty()) die("--obj and --camera-sig are required");

    /* Renderer::setMesh / getCudaBVH -- Renderer.cc:53-54,77-99,157-217 */
    MeshBase* mesh = importMesh(String(obj.c_str()));
    if (!mesh) die("importMesh failed");
    Scene scene(*mesh);
    Platform platform("GPU");
    platform.setLeafPreferences(1, 8);
    BVH::BuildParams params;
    params.enablePrints = false;
    BVH::Stats stats;
    params.stats = &stats;
    BVH bvh(&scene, platform, params);
    CudaBVH cbvh(bvh, BVHLayout_Compact2);
    {
        FileWR f(String((out + ".bvh").c_str()));
        cbvh.serialize(f);
    }
    fprintf(stderr, "route_a: %d tris, %d inner nodes, %d Woop rows\n", scene.getNumTriangles(), stats.numInnerNodes,
            (int)(cbvh.getTriWoopBuffer().getSize() / 16));
    if (cbvh.getTriWoopBuffer().getSize() / 16 < 750000) die("scene too small for CudaTracer.cc's hard-coded triangle dump (750 000 rows)");

    /* Renderer::beginFrame -- Renderer.cc:112-152 */
    CameraControls camera;
    camera.decodeSignature(String(camSig.c_str()));
    Mat4f nscreenToWorld = invert(Mat4f::fitToView(-1.0f, 2.0f, Vec2i(W, H)) * camera.getWorldToClip());
    CudaTracer tracer;
    tracer.initKernel();
    tracer.second_fetch = false;
    tracer.setBVH(&cbvh);
    RayGen raygen(1 << 21);
    RayBuffer primary;
    raygen.primary(primary, camera.getPosition(), nscreenToWorld, W, H, camera.getFar());
    F32 msPrimary = tracer.traceBatch(primary);
    dump(out + ".prim.rays", primary.getRayBuffer_dev(), primary.getSize(), sizeof(Ray));
    dump(out + ".prim.results", primary.getResultBuffer_dev(), primary.getSize(), sizeof(RayResult));

    /* Renderer::nextBatch, Diffuse with sortSecondary -- Renderer.cc:242-291 */
    RayBuffer secondary;
    bool newBatch = true;
    if (!raygen.ao(secondary, primary, scene, samples, camera.getFar(), newBatch, 0)) die("RayGen::ao produced no batch");
    secondary.setNeedClosestHit(true);
    dump(out + ".sec.unsorted.rays", secondary.getRayBuffer_dev(), secondary.getSize(), sizeof(Ray));
    secondary.mortonSort();
    dump(out + ".sec.rays", secondary.getRayBuffer_dev(), secondary.getSize(), sizeof(Ray));
    dump(out + ".sec.slot2id", secondary.getSlotToIDBuffer_dev(), secondary.getSize(), 4);
    dump(out + ".sec.id2slot", secondary.getIDToSlotBuffer_dev(), secondary.getSize(), 4);
    F32 msSecondary = tracer.traceBatch(secondary);
    dump(out + ".sec.results", secondary.getResultBuffer_dev(), secondary.getSize(), sizeof(RayResult));
    fprintf(stdout, "{\"primary_rays\": %d, \"primary_ms\": %.4f, \"secondary_rays\": %d, \"secondary_ms\": %.4f}\n",
            primary.getSize(), msPrimary, secondary.getSize(), msSecondary);
    return 0;
}
