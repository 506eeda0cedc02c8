/*
 * ref_driver.cc -- command-line driver around the UNMODIFIED reference host code.
 *
 * TEST INFRASTRUCTURE ONLY (oracle/_ref/ref_cpu).  This file is ours; it includes the
 * reference's headers and is linked against the reference's own .cc files compiled from
 * where they lie under /root/reference (recipe: oracle/Makefile).  No reference source is
 * copied into the repository.
 *
 * What runs reference code: OBJ import (Mesh.cc / MeshWavefrontIO.cc), Scene, SplitBVHBuilder,
 * CudaBVH::createCompact + woopifyTri, CudaBVH::serialize (bvhcache format), CameraControls,
 * PixelTable, RayGen::primary / RayGen::ao host logic (batching, rand() seed), RayBuffer
 * (id<->slot maps, mortonSort with FW::sort), BVH::trace.
 * The reference's ray-generation, Morton-key, reorder and reconstruct KERNELS run too: their .cu
 * files are compiled as host C++ with gpu_ref/host_emul.h and every emulated thread executes the
 * reference's own kernel body (launch_rayGenPrimaryKernel ... launch_reorderRaysKernel,
 * RayGenKernels.hh:107-111, RayBufferKernels.hh:112-117, RendererKernels.hh:88-92).
 * What is ours: launch_findAABBKernel (a block-cooperative reduction a serial loop cannot emulate)
 * is bound to b2rt_oracle.c; cudaMalloc/cudaMemcpy/cudaFree are bound to malloc/memcpy/free
 * (stubs.c), so "device" pointers are host pointers.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <chrono>
#include <unistd.h>
#include <sys/wait.h>
#include <sys/mman.h>

#include "base/Defs.hh"
#include "base/Math.hh"
#include "3d/Mesh.hh"
#include "3d/CameraControls.hh"
#include "io/File.hh"
#include "Scene.hh"
#include "bvh/BVH.hh"
#include "cuda/CudaBVH.hh"
#include "ray/RayBuffer.hh"
#include "ray/RayGen.hh"
#include "ray/RayGenKernels.hh"
#include "ray/RayBufferKernels.hh"

#include "b2rt_oracle.h"

using namespace FW;

/* ---- launchers the reference host code calls ----
 * launch_rayGenPrimaryKernel / launch_rayGenAOKernel / launch_genMortonKeysKernel /
 * launch_reorderRaysKernel / launch_reconstructKernel are the REFERENCE'S OWN (RayGenKernels.cu,
 * RayBufferKernels.cu, RendererKernels.cu compiled as host C++, kernel bodies run serially: see
 * gpu_ref/host_emul.h).  Only the two block-cooperative reductions are bound to the restatement. */
extern "C" void launch_findAABBKernel(S32, const Vec2i&, struct FindAABBInput& in, struct FindAABBOutput& out)
{
    b2o_find_aabb((const b2o_ray*)in.inRays, in.numRays, out.aabbLo.getPtr(), out.aabbHi.getPtr());
}

/* ---- helpers ---- */
static void die(const char* msg) { fprintf(stderr, "ref_cpu: %s\n", msg); exit(2); }

static void writeFile(const std::string& path, const void* a, size_t na, const void* b = NULL, size_t nb = 0,
                      const void* c = NULL, size_t nc = 0, const void* d = NULL, size_t nd = 0)
{
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) die(("cannot write " + path).c_str());
    if (na) fwrite(a, 1, na, f);
    if (nb) fwrite(b, 1, nb, f);
    if (nc) fwrite(c, 1, nc, f);
    if (nd) fwrite(d, 1, nd, f);
    fclose(f);
}

/* rays file = the reference's dumpRayBuffer format (RayBuffer.cc:89-140): int32 count, N x 32 B */
static void dumpRays(const std::string& path, const Ray* rays, S32 n) { writeFile(path, &n, 4, rays, (size_t)n * sizeof(Ray)); }
static void dumpInts(const std::string& path, const S32* v, S32 n) { writeFile(path, &n, 4, v, (size_t)n * 4); }
static void dumpResults(const std::string& path, const RayResult* r, S32 n) { writeFile(path, &n, 4, r, (size_t)n * sizeof(RayResult)); }

static void devToHost(RayBuffer& rb)
{
    S32 n = rb.getSize();
    if (!n) return;
    memcpy(rb.getRayBuffer().getMutablePtr(), rb.getRayBuffer_dev(), (size_t)n * sizeof(Ray));
    memcpy(rb.getIDToSlotBuffer().getMutablePtr(), rb.getIDToSlotBuffer_dev(), (size_t)n * 4);
    memcpy(rb.getSlotToIDBuffer().getMutablePtr(), rb.getSlotToIDBuffer_dev(), (size_t)n * 4);
}
static void resultsHostToDev(RayBuffer& rb)
{
    if (rb.getSize())
        memcpy(rb.getResultBuffer_dev(), rb.getResultBuffer().getPtr(), (size_t)rb.getSize() * sizeof(RayResult));
}

/* BVH::trace over [0,n) split across `procs` forked workers (BVH.cc:64-65 keeps per-ray state
 * in file-scope statics, so threads would race; processes do not).  Returns wall seconds. */
static double traceTimed(const BVH& bvh, RayBuffer& rb, int procs)
{
    S32 n = rb.getSize();
    auto t0 = std::chrono::steady_clock::now();
    if (procs <= 1) {
        bvh.trace(rb);
    } else {
        RayResult* shared = (RayResult*)mmap(NULL, (size_t)n * sizeof(RayResult), PROT_READ | PROT_WRITE,
                                             MAP_SHARED | MAP_ANONYMOUS, -1, 0);
        if (shared == MAP_FAILED) die("mmap failed");
        std::vector<pid_t> kids;
        for (int p = 0; p < procs; p++) {
            S32 lo = (S32)((S64)n * p / procs), hi = (S32)((S64)n * (p + 1) / procs);
            pid_t pid = fork();
            if (pid < 0) die("fork failed");
            if (pid == 0) {
                RayBuffer part(hi - lo, rb.getNeedClosestHit());
                for (S32 i = lo; i < hi; i++) part.setRay(i - lo, rb.getRayForSlot(i));
                bvh.trace(part);
                memcpy(shared + lo, part.getResultBuffer().getPtr(), (size_t)(hi - lo) * sizeof(RayResult));
                _exit(0);
            }
            kids.push_back(pid);
        }
        for (pid_t k : kids) { int st; waitpid(k, &st, 0); if (!WIFEXITED(st) || WEXITSTATUS(st)) die("worker failed"); }
        memcpy(rb.getResultBuffer().getMutablePtr(), shared, (size_t)n * sizeof(RayResult));
        munmap(shared, (size_t)n * sizeof(RayResult));
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

/* Bounded sample: every `stride`-th run of 64 consecutive slots (one 8x8 pixel block of the
 * PixelTable order, PixelTable.cc:94-156), so the sample covers the whole frame evenly. */
static void strideSample(RayBuffer& rb, int stride)
{
    if (stride <= 1) return;
    S32 n = rb.getSize();
    std::vector<Ray> keep;
    for (S32 b = 0; b * 64 < n; b += stride)
        for (S32 i = b * 64; i < n && i < b * 64 + 64; i++) keep.push_back(rb.getRayForSlot(i));
    bool closest = rb.getNeedClosestHit();
    rb.resize((S32)keep.size());
    rb.setNeedClosestHit(closest);
    for (S32 i = 0; i < (S32)keep.size(); i++) rb.setRay(i, keep[i]);
    memcpy(rb.getRayBuffer_dev(), keep.data(), keep.size() * sizeof(Ray));
}

static void usage(void)
{
    fprintf(stderr,
        "ref_cpu --obj F [--alpha A] [--scene-out F] [--bvh-out F]\n"
        "        [--camera-sig S | --camera-auto] [--size W H] [--camera-out F]\n"
        "        [--primary-out PREFIX]            rays via RayGen::primary -> PREFIX.rays/.slot2id\n"
        "        [--rays-in F [--any-hit]]         trace these instead of the primaries\n"
        "        [--results-out F] [--procs P] [--limit N] [--repeat K] [--budget-s T | --prim-stride S]   BVH::trace, timed (one JSON line per repeat)\n"
        "        [--go-file F]                     after the BVH build, wait until file F exists before tracing\n"
        "        [--ao-out PREFIX --samples S --max-dist D [--sort]]   RayGen::ao batches from the primaries\n"
        "        [--trace-secondary [--secondary-closest] [--secondary-results-out F]]  time BVH::trace on the first secondary batch\n"
        "        [--sort-in F --sort-out PREFIX]   RayBuffer::mortonSort\n"
        "        [--cache-key]                     the bvhcache file name key, computed by the reference's own hash code\n");
    exit(2);
}

int main(int argc, char** argv)
{
    std::string obj, sceneOut, bvhOut, camSig, camOut, primOut, raysIn, resOut, aoOut, sortIn, sortOut;
    bool camAuto = false, anyHit = false, aoSort = false, secTrace = false, secClosest = false;
    std::string secResOut;
    bool cacheKey = false;
    float alpha = 1.0e-5f, maxDist = 0.0f;
    int W = 1024, H = 768, procs = 1, samples = 1, limit = -1, repeat = 1;
    double budgetS = 0.0;
    int primStride = 0;
    std::string goFile;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { if (i + 1 >= argc) usage(); return argv[++i]; };
        if (a == "--obj") obj = next();
        else if (a == "--alpha") alpha = (float)atof(next());
        else if (a == "--scene-out") sceneOut = next();
        else if (a == "--bvh-out") bvhOut = next();
        else if (a == "--camera-sig") camSig = next();
        else if (a == "--camera-auto") camAuto = true;
        else if (a == "--size") { W = atoi(next()); H = atoi(next()); }
        else if (a == "--camera-out") camOut = next();
        else if (a == "--primary-out") primOut = next();
        else if (a == "--rays-in") raysIn = next();
        else if (a == "--any-hit") anyHit = true;
        else if (a == "--results-out") resOut = next();
        else if (a == "--procs") procs = atoi(next());
        else if (a == "--limit") limit = atoi(next());
        else if (a == "--repeat") repeat = atoi(next());
        else if (a == "--budget-s") budgetS = atof(next());      /* bound the total traced time by sampling the frame */
        else if (a == "--prim-stride") primStride = atoi(next()); /* keep every S-th 8x8 pixel block of the frame (explicit form of --budget-s) */
        else if (a == "--go-file") goFile = next();               /* lets a caller overlap the single-threaded BVH build with its own work */
        else if (a == "--ao-out") aoOut = next();
        else if (a == "--samples") samples = atoi(next());
        else if (a == "--max-dist") maxDist = (float)atof(next());
        else if (a == "--sort") aoSort = true;
        else if (a == "--trace-secondary") secTrace = true;       /* time BVH::trace on the first secondary batch */
        else if (a == "--secondary-closest") secClosest = true;   /* Diffuse: setNeedClosestHit(true), Renderer.cc:277 */
        else if (a == "--secondary-results-out") secResOut = next();
        else if (a == "--cache-key") cacheKey = true;             /* print Renderer::getCudaBVH's cache file key (Renderer.cc:180-186) */
        else if (a == "--sort-in") sortIn = next();
        else if (a == "--sort-out") sortOut = next();
        else usage();
    }

    /* mortonSort alone needs no scene */
    if (!sortIn.empty()) {
        FILE* f = fopen(sortIn.c_str(), "rb"); if (!f) die("cannot read --sort-in");
        S32 n = 0; if (fread(&n, 4, 1, f) != 1) die("short rays file");
        RayBuffer rb(n, true);
        std::vector<Ray> tmp(n);
        if (n && fread(tmp.data(), sizeof(Ray), n, f) != (size_t)n) die("short rays file");
        fclose(f);
        memcpy(rb.getRayBuffer_dev(), tmp.data(), (size_t)n * sizeof(Ray));
        for (S32 i = 0; i < n; i++) { rb.getIDToSlotBuffer_dev()[i] = i; rb.getSlotToIDBuffer_dev()[i] = i; }
        rb.mortonSort();
        dumpRays(sortOut + ".rays", rb.getRayBuffer_dev(), n);
        dumpInts(sortOut + ".slot2id", rb.getSlotToIDBuffer_dev(), n);
        dumpInts(sortOut + ".id2slot", rb.getIDToSlotBuffer_dev(), n);
        Vec3f lo, hi; rb.getRayAABB(lo, hi);
        float box[6] = { lo.x, lo.y, lo.z, hi.x, hi.y, hi.z };
        writeFile(sortOut + ".aabb", box, sizeof(box));
        if (obj.empty()) return 0;
    }
    if (obj.empty()) usage();

    /* Renderer::setMesh / getCudaBVH -- Renderer.cc:53-54,77-99,157-217 */
    MeshBase* mesh = importMesh(String(obj.c_str()));
    if (!mesh) die("importMesh failed");
    Scene scene(*mesh);
    Platform platform("GPU");
    platform.setLeafPreferences(1, 8);
    BVH::BuildParams params;
    params.splitAlpha = alpha;
    params.enablePrints = false;
    BVH::Stats stats;
    params.stats = &stats;
    auto tb0 = std::chrono::steady_clock::now();
    BVH bvh(&scene, platform, params);
    double buildSec = std::chrono::duration<double>(std::chrono::steady_clock::now() - tb0).count();
    CudaBVH cbvh(bvh, BVHLayout_Compact2);
    fprintf(stderr, "ref_cpu: %d tris %d verts; BVH inner %d leaf %d refs %d sah %.3f build %.2fs\n",
            scene.getNumTriangles(), scene.getNumVertices(), stats.numInnerNodes, stats.numLeafNodes,
            stats.numTris, stats.SAHCost, buildSec);

    if (cacheKey) {
        /* exactly the expression of Renderer::getCudaBVH (Renderer.cc:180-186); the pieces: Scene.cc:94-102, Platform.hh:69, BVH.hh:82-85 */
        U32 hs = scene.hash(), hp = platform.computeHash(), hb = params.computeHash();
        fprintf(stdout, "{\"cache_key\": \"%08x\", \"scene_hash\": %u, \"platform_hash\": %u, \"params_hash\": %u, \"layout\": %d}\n",
                hashBits(hs, hp, hb, (U32)BVHLayout_Compact2), hs, hp, hb, (int)BVHLayout_Compact2);
        fflush(stdout);
    }
    if (!sceneOut.empty()) {
        S32 hdr[3] = { 0x43533242 /* "B2SC" */, scene.getNumVertices(), scene.getNumTriangles() };
        FILE* f = fopen(sceneOut.c_str(), "wb"); if (!f) die("cannot write --scene-out");
        fwrite(hdr, 4, 3, f);
        fwrite(scene.getVtxPosBuffer().getPtr(), 12, hdr[1], f);
        fwrite(scene.getTriVtxIndexBuffer().getPtr(), 12, hdr[2], f);
        fwrite(scene.getTriNormalBuffer().getPtr(), 12, hdr[2], f);
        fclose(f);
    }
    if (!bvhOut.empty()) {
        FileWR out(String(bvhOut.c_str()));
        cbvh.serialize(out);      /* bvhcache format: S32 layout + 3 x (S64 size, bytes) */
    }

    /* camera -- App.cc:174, Renderer.cc:126-129 */
    CameraControls camera;
    if (!camSig.empty()) camera.decodeSignature(String(camSig.c_str()));
    else if (camAuto) camera.initForMesh(mesh);
    Mat4f nscreenToWorld = invert(Mat4f::fitToView(-1.0f, 2.0f, Vec2i(W, H)) * camera.getWorldToClip());
    if (!camOut.empty()) {
        float buf[20];
        memcpy(buf, camera.getPosition().getPtr(), 12);
        memcpy(buf + 3, nscreenToWorld.getPtr(), 64);
        buf[19] = camera.getFar();
        writeFile(camOut, buf, sizeof(buf));
    }

    if (!goFile.empty()) {
        fprintf(stderr, "ref_cpu: ready, waiting for %s\n", goFile.c_str());
        while (access(goFile.c_str(), F_OK) != 0) usleep(20000);
    }
    RayGen raygen(1 << 21);       /* Renderer.cc:46 */
    RayBuffer primary;
    bool havePrimary = false;
    if (!raysIn.empty()) {
        FILE* f = fopen(raysIn.c_str(), "rb"); if (!f) die("cannot read --rays-in");
        S32 n = 0; if (fread(&n, 4, 1, f) != 1) die("short rays file");
        if (limit >= 0 && limit < n) n = limit;
        primary.resize(n);
        primary.setNeedClosestHit(!anyHit);
        std::vector<Ray> tmp(n);
        if (n && fread(tmp.data(), sizeof(Ray), n, f) != (size_t)n) die("short rays file");
        fclose(f);
        for (S32 i = 0; i < n; i++) primary.setRay(i, tmp[i]);
        memcpy(primary.getRayBuffer_dev(), tmp.data(), (size_t)n * sizeof(Ray));
        havePrimary = true;
    } else if (!primOut.empty() || !resOut.empty() || !aoOut.empty() || secTrace) {
        raygen.primary(primary, camera.getPosition(), nscreenToWorld, W, H, camera.getFar());
        devToHost(primary);
        havePrimary = true;
        if (!primOut.empty()) {
            dumpRays(primOut + ".rays", (const Ray*)primary.getRayBuffer().getPtr(), primary.getSize());
            dumpInts(primOut + ".slot2id", (const S32*)primary.getSlotToIDBuffer().getPtr(), primary.getSize());
        }
    }

    if (havePrimary && (!resOut.empty() || !aoOut.empty() || secTrace)) {
        if (primStride > 1 && raysIn.empty()) {
            strideSample(primary, primStride);
            fprintf(stderr, "ref_cpu: every %d-th 8x8 block -> %d primary rays\n", primStride, primary.getSize());
        } else
        if (budgetS > 0.0 && raysIn.empty()) {
            /* calibrate on a thin sample, then keep every stride-th 8x8 block of the frame */
            RayBuffer probe(primary.getSize(), true);
            for (S32 i = 0; i < primary.getSize(); i++) probe.setRay(i, primary.getRayForSlot(i));
            int pstride = primary.getSize() / 64 / 1024; if (pstride < 1) pstride = 1;
            strideSample(probe, pstride);
            double ps = traceTimed(bvh, probe, procs);
            double rate = probe.getSize() / ps;
            double want = budgetS * rate / ((double)repeat * (secTrace ? 2 : 1));
            int stride = (int)((double)primary.getSize() / want + 0.999999);
            if (stride < 1) stride = 1;
            strideSample(primary, stride);
            fprintf(stderr, "ref_cpu: budget %.0fs, probe %.3f Mrays/s -> stride %d, %d rays per batch\n",
                    budgetS, rate * 1e-6, stride, primary.getSize());
        }
        S32 n = primary.getSize();
        double sec = 0.0;
        for (int rep = 0; rep < repeat; rep++) {
        sec = traceTimed(bvh, primary, procs);
        fprintf(stdout, "{\"rays\": %d, \"trace_s\": %.6f, \"mrays_per_s\": %.4f, \"procs\": %d, \"closest\": %d}\n",
                n, sec, n / sec * 1e-6, procs, primary.getNeedClosestHit() ? 1 : 0);
        fflush(stdout);
        }
        if (!resOut.empty())
            dumpResults(resOut, (const RayResult*)primary.getResultBuffer().getPtr(), n);
    }

    /* Renderer::nextBatch -- Renderer.cc:242-291: AO / diffuse batches from the primary hits */
    if (secTrace) {
        resultsHostToDev(primary);
        RayBuffer secondary;
        bool newBatch = true;
        if (raygen.ao(secondary, primary, scene, samples, maxDist, newBatch, 0)) {
            if (secClosest) secondary.setNeedClosestHit(true);
            if (aoSort) secondary.mortonSort();
            if (limit >= 0 && limit < secondary.getSize()) secondary.resize(limit);
            devToHost(secondary);
            S32 n = secondary.getSize(), counted = 0;
            for (S32 i = 0; i < n; i++) if (secondary.getRayForSlot(i).tmax >= 0.0f) counted++;
            for (int rep = 0; rep < repeat; rep++) {
                double sec = traceTimed(bvh, secondary, procs);
                fprintf(stdout, "{\"stage\": \"secondary\", \"rays\": %d, \"counted\": %d, \"trace_s\": %.6f, \"mrays_per_s\": %.4f, \"procs\": %d, \"closest\": %d}\n",
                        n, counted, sec, counted / sec * 1e-6, procs, secondary.getNeedClosestHit() ? 1 : 0);
                fflush(stdout);
            }
            if (!secResOut.empty()) dumpResults(secResOut, (const RayResult*)secondary.getResultBuffer().getPtr(), n);
            if (!aoOut.empty()) dumpRays(aoOut + ".rays", (const Ray*)secondary.getRayBuffer().getPtr(), n);
        }
    } else if (!aoOut.empty()) {
        resultsHostToDev(primary);
        RayBuffer secondary;
        bool newBatch = true;
        int batch = 0;
        while (raygen.ao(secondary, primary, scene, samples, maxDist, newBatch, 0)) {
            if (aoSort) secondary.mortonSort();
            char suffix[64]; snprintf(suffix, sizeof(suffix), ".%02d", batch++);
            dumpRays(aoOut + suffix + ".rays", secondary.getRayBuffer_dev(), secondary.getSize());
            dumpInts(aoOut + suffix + ".slot2id", secondary.getSlotToIDBuffer_dev(), secondary.getSize());
        }
        fprintf(stderr, "ref_cpu: %d secondary batches\n", batch);
    }
    return 0;
}
