"""Generates the golden fixtures of tests/golden/ by running the REFERENCE'S OWN code
(oracle/_ref/ref_cpu: the reference's host sources and, through gpu_ref/host_emul.h, its ray
generation / Morton key / reorder kernel bodies, compiled from /root/reference by oracle/Makefile).

    python tests/golden/make_golden.py        (needs /root/reference; run in the build container)

The fixtures pin the oracle (oracle/b2rt_oracle.c) on machines where the reference is absent:
tests/test_oracle_golden.py checks every b2o_* function against them.  Inputs are two small
seeded scenes written as OBJ; everything else in the .npz files is reference output.
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as O      # noqa: E402

W, H = 40, 24          # not multiples of 8 in x: exercises PixelTable's remainder stripes (PixelTable.cc:128-156)
SAMPLES = 4


def scene_blob(seed, n_lat=10, n_lon=14):
    """Bumpy sphere: closed mesh, shared edges and vertices everywhere."""
    rng = np.random.RandomState(seed)
    verts = [[0.0, 1.0, 0.0]]
    for i in range(1, n_lat):
        th = np.pi * i / n_lat
        for j in range(n_lon):
            ph = 2 * np.pi * j / n_lon
            r = 1.0 + 0.15 * rng.rand()
            verts.append([r * np.sin(th) * np.cos(ph), r * np.cos(th), r * np.sin(th) * np.sin(ph)])
    verts.append([0.0, -1.0, 0.0])
    tris = []
    for j in range(n_lon):
        tris.append([0, 1 + j, 1 + (j + 1) % n_lon])
    for i in range(n_lat - 2):
        a, b = 1 + i * n_lon, 1 + (i + 1) * n_lon
        for j in range(n_lon):
            j2 = (j + 1) % n_lon
            tris.append([a + j, b + j, b + j2])
            tris.append([a + j, b + j2, a + j2])
    last = len(verts) - 1
    a = 1 + (n_lat - 2) * n_lon
    for j in range(n_lon):
        tris.append([last, a + (j + 1) % n_lon, a + j])
    return np.asarray(verts, np.float32), np.asarray(tris, np.int32)


def scene_room(seed):
    """Box room seen from inside with a few boxes in it: large + small triangles, axis-aligned
    geometry (rays with zero direction components, coplanar shared edges)."""
    rng = np.random.RandomState(seed)
    verts, tris = [], []

    def box(lo, hi, inward):
        base = len(verts)
        for z in (lo[2], hi[2]):
            for y in (lo[1], hi[1]):
                for x in (lo[0], hi[0]):
                    verts.append([x, y, z])
        quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
        for q in quads:
            q = q[::-1] if inward else q
            tris.append([base + q[0], base + q[1], base + q[2]])
            tris.append([base + q[0], base + q[2], base + q[3]])

    box((-4, 0, -3), (4, 3, 3), True)
    for _ in range(12):
        c = np.array([rng.uniform(-3, 3), 0.0, rng.uniform(-2.2, 2.2)])
        s = rng.uniform(0.15, 0.6, 3)
        box(c - [s[0], 0, s[2]], c + [s[0], 2 * s[1], s[2]], False)
    return np.asarray(verts, np.float32), np.asarray(tris, np.int32)


def run_case(name, verts, tris, camera_args, tmp):
    obj = os.path.join(HERE, name + ".obj")
    O.write_obj(obj, verts, tris)
    p = lambda s: os.path.join(tmp, name + "_" + s)
    # scene, BVH, camera, primary rays, BVH::trace of the primaries
    O.run_ref_cpu("--obj", obj, "--scene-out", p("scene"), "--bvh-out", p("bvh"), *camera_args, "--size", W, H,
                  "--camera-out", p("cam"), "--primary-out", p("prim"), "--results-out", p("prim.results"))
    scene, bvh, cam = O.read_scene(p("scene")), O.read_bvh(p("bvh")), O.read_camera(p("cam"))
    prim, prim_s2i, prim_res = O.read_rays(p("prim.rays")), O.read_ints(p("prim.slot2id")), O.read_results(p("prim.results"))
    # AO batch (any hit), Morton-sorted, and a diffuse batch (closest hit) with their BVH::trace results
    diag = float(np.linalg.norm(verts.max(0) - verts.min(0)))
    out = {}
    for tag, extra, maxdist in (("ao", ["--sort"], 0.3 * diag), ("diffuse", ["--secondary-closest"], cam["far"])):
        O.run_ref_cpu("--obj", obj, *camera_args, "--size", W, H, "--trace-secondary", *extra, "--samples", SAMPLES,
                      "--max-dist", maxdist, "--ao-out", p(tag), "--secondary-results-out", p(tag + ".results"))
        out[tag + "_rays"] = O.read_rays(p(tag + ".rays"))
        out[tag + "_results"] = O.read_results(p(tag + ".results"))
        out[tag + "_maxdist"] = np.float32(maxdist)
    # unsorted AO batch + ids (RayGen::ao straight out of the reference kernel body)
    O.run_ref_cpu("--obj", obj, *camera_args, "--size", W, H, "--samples", SAMPLES, "--max-dist", 0.3 * diag, "--ao-out", p("aoraw"))
    out["aoraw_rays"] = O.read_rays(p("aoraw.00.rays"))
    out["aoraw_slot2id"] = O.read_ints(p("aoraw.00.slot2id"))
    # RayBuffer::mortonSort on the unsorted batch (reference key kernel + FW::sort + reorder kernel)
    O.write_rays(p("sortin.rays"), out["aoraw_rays"])
    O.run_ref_cpu("--sort-in", p("sortin.rays"), "--sort-out", p("sorted"))
    out["sorted_rays"] = O.read_rays(p("sorted.rays"))
    out["sorted_slot2id"] = O.read_ints(p("sorted.slot2id"))
    out["sorted_id2slot"] = O.read_ints(p("sorted.id2slot"))
    out["sorted_aabb"] = np.fromfile(p("sorted.aabb"), np.float32)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), verts=scene["verts"], tris=scene["tris"], normals=scene["normals"],
                        nodes=bvh["nodes"], woop=bvh["woop"], triIndex=bvh["triIndex"], layout=np.int32(bvh["layout"]),
                        cam_origin=cam["origin"], cam_nscreenToWorld=cam["nscreenToWorld"], cam_far=np.float32(cam["far"]),
                        size=np.array([W, H], np.int32), samples=np.int32(SAMPLES), ao_seed=np.uint32(1804289383),
                        prim_rays=prim, prim_slot2id=prim_s2i, prim_results=prim_res, **out)
    print(name, "tris", len(tris), "nodes", bvh["nodes"].size // 16, "prim hits", int((prim_res["id"] >= 0).sum()), "/", len(prim))


def main():
    if not O.have_ref_cpu():
        O.build_oracle()
    assert O.have_ref_cpu(), "oracle/_ref/ref_cpu missing: needs /root/reference"
    tmp = "/tmp/b2rt_golden"
    os.makedirs(tmp, exist_ok=True)
    v, t = scene_blob(1)
    run_case("blob", v, t, ["--camera-auto"], tmp)
    v, t = scene_room(2)
    # camera inside the room, looking down +x/-z, as a --camera= signature (CameraControls.cc:354-419)
    import importlib
    sys.path.insert(0, ROOT)
    host = importlib.import_module("gpu-ray-tracing_b200.host")
    cam = {"pos": [-3.2, 1.6, 2.4], "forward": [0.83, -0.2, -0.52], "up": [0, 1, 0], "fov": 70.0, "near": 0.01, "far": 30.0}
    sig = host.encode_camera_signature(cam).strip('",')
    with open(os.path.join(HERE, "room.camera.txt"), "w") as f:
        f.write(sig + "\n")
    run_case("room", v, t, ["--camera-sig", sig], tmp)


if __name__ == "__main__":
    main()
