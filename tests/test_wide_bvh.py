"""Host collapse to the private 8-wide compressed BVH (csrc/host/wide_bvh.cpp), checked on the CPU:
structure, and a scalar Python traversal of the wide form against the oracle's traversal of the
binary Compact2 arrays on the same rays (closest hit is order-independent, so ids and t must agree)."""
import os

import numpy as np
import pytest

import wide_ref

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _check(b2rt, oracle, arrays, rays, min_agree=0.995):
    wide = wide_ref.build(b2rt, arrays)
    ref = oracle.trace_woop(arrays, rays)
    same = 0
    visits = 0
    for i, r in enumerate(rays):
        hit, t, nn, nt = wide_ref.trace(wide, r)
        visits += nn
        rid = -1 if hit < 0 else int(wide["triIndex"][hit])
        if rid == ref["id"][i]:
            same += 1
            if rid >= 0:
                assert abs(t - ref["t"][i]) <= 1e-5 * abs(ref["t"][i]) + 1e-7
        else:
            # only ties between two triangles at the same distance may differ
            assert rid >= 0 and ref["id"][i] >= 0 and abs(t - ref["t"][i]) <= 1e-4 * abs(ref["t"][i]), "ray %d" % i
    assert same >= min_agree * len(rays)
    return wide, visits / len(rays)


@pytest.mark.parametrize("case", ["blob", "room"])
def test_wide_on_reference_built_bvh(b2rt, oracle, case):
    """Collapse of a BVH built by the REFERENCE's SplitBVHBuilder + createCompact (golden arrays)."""
    g = dict(np.load(os.path.join(GOLDEN, case + ".npz")))
    arrays = {"nodes": g["nodes"], "woop": g["woop"], "triIndex": g["triIndex"]}
    rays = np.concatenate([g["prim_rays"][::3], g["diffuse_rays"][::16]])
    wide, visits = _check(b2rt, oracle, arrays, rays)
    assert wide["stats"]["nodes"] < g["nodes"].size // 16          # collapsed
    # every triangle reference of the binary tree is kept exactly once
    # (walk the rows: only the FIRST row of a triangle is guaranteed not to look like a terminator, CudaBVH.cc:316-317)
    bits = g["woop"].reshape(-1, 4)[:, 0].view(np.uint32)
    n_refs, i = 0, 0
    while i < len(bits):
        if bits[i] == 0x80000000:
            i += 1
        else:
            n_refs += 1
            i += 3
    assert wide["stats"]["tris"] == n_refs


@pytest.mark.parametrize("kind", ["interior", "blob"])
def test_wide_on_product_built_bvh(b2rt, oracle, kind):
    scene = b2rt.Scene.generate(kind, 3000, 17)
    bvh = b2rt.BVH(scene)
    origin, m = b2rt.camera_matrices(scene.camera, 20, 12)
    i2p, _ = oracle.pixel_table(20, 12)
    rays, _, _ = oracle.raygen_primary(origin, m, 20, 12, scene.camera["far"], i2p)
    rays = rays.copy()
    if kind == "blob":
        # axis-parallel components.  Not in the axis-aligned interior: there such rays lie exactly IN the planes of box
        # faces and triangle edges, where the binary tree's exact boxes cull a leaf that the (slightly larger) quantised
        # boxes keep -- both answers are legitimate for a hit exactly on an edge, but they are not the same id.
        rays["d"][::7, 1] = 0.0
    rays["tmax"][::11] = -1.0                                       # degenerate rays must miss
    wide, visits = _check(b2rt, oracle, bvh.arrays(), rays)
    bin_visits = oracle.trace_woop(bvh.arrays(), rays, counters=True)[1]["nodes"] / len(rays)
    assert visits < 0.6 * bin_visits                                # the point of going wide


def test_wide_build_rejects_bad_input(b2rt):
    import ctypes as C
    h = C.c_void_p()
    bad = np.zeros(16, np.float32)
    bad[12:14] = np.array([400, 800], np.int32).view(np.float32)    # children beyond the array
    w = np.zeros(4, np.float32)
    t = np.zeros(1, np.int32)
    rc = b2rt.lib().b2rt_wide_build(bad.ctypes.data_as(C.c_void_p), bad.nbytes, w.ctypes.data_as(C.c_void_p), w.nbytes,
                                    t.ctypes.data_as(C.c_void_p), t.nbytes, C.byref(h))
    assert rc != 0 and not h.value
