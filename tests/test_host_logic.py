"""Host-side logic of the product (no GPU): pixel table, camera, scene generators, SBVH builder and
Compact2 flattening, bvhcache file format, batching, sharding -- each against the oracle."""
import os
import tempfile

import numpy as np
import pytest


@pytest.mark.parametrize("w,h", [(8, 8), (40, 24), (203, 77), (5, 3), (1024, 768)])
def test_pixel_table_matches_oracle(b2rt, oracle, w, h):
    pt = b2rt.PixelTable()
    i2p = np.empty(w * h, np.int32)
    p2i = np.empty(w * h, np.int32)
    b2rt._lib.check(b2rt.lib().b2rt_pixel_table(w, h, b2rt._lib.np_ptr(i2p), b2rt._lib.np_ptr(p2i)))
    ri2p, rp2i = oracle.pixel_table(w, h)
    assert (i2p == ri2p).all() and (p2i == rp2i).all()
    assert sorted(i2p.tolist()) == list(range(w * h))


def test_camera_signature_round_trip(b2rt):
    cam = {"pos": np.array([1.25, -2.5, 7.0], np.float32), "forward": np.array([0.6, -0.48, 0.64], np.float32),
           "up": np.array([0, 1, 0], np.float32), "fov": 55.0, "near": 0.02, "far": 123.0}
    sig = b2rt.encode_camera_signature(cam)
    back = b2rt.decode_camera_signature(sig)
    assert np.allclose(back["pos"], cam["pos"]) and np.allclose(back["forward"], cam["forward"], atol=1e-6)
    assert back["fov"] == 55.0 and abs(back["far"] - 123.0) < 1e-4
    with pytest.raises(RuntimeError, match="Invalid signature"):
        b2rt.decode_camera_signature('"!!!"')


def test_camera_matrices_match_oracle(b2rt, oracle):
    cam = {"pos": [0.5, 1.0, 4.0], "forward": [0.0, -0.2, -1.0], "up": [0, 1, 0], "fov": 70.0, "near": 0.01, "far": 50.0}
    o, m = b2rt.camera_matrices(cam, 1024, 768)
    ro, rm = oracle.camera_matrices(cam["pos"], cam["forward"], cam["up"], 70.0, 0.01, 50.0, 1024, 768)
    assert (o == ro).all() and (m == rm).all()


@pytest.mark.parametrize("kind,target", [("blob", 3000), ("interior", 5000), ("knot", 4000), ("strands", 6000)])
def test_scene_generators(b2rt, oracle, kind, target):
    a = b2rt.Scene.generate(kind, target, 42)
    b = b2rt.Scene.generate(kind, target, 42)
    assert abs(a.getNumTriangles() - target) <= 0.005 * target + 8
    assert a.verts.tobytes() == b.verts.tobytes() and a.tris.tobytes() == b.tris.tobytes()      # seeded
    assert a.tris.min() >= 0 and a.tris.max() < a.getNumVertices()
    assert (a.normals == oracle.tri_normals(a.verts, a.tris)).all()                             # Scene.cc:72
    assert a.camera["far"] > a.camera["near"] > 0


def _brute_force(scene, rays):
    """Closest hit over ALL triangles with the oracle's Moller-Trumbore (a one-leaf BVH)."""
    ids = np.full(len(rays), -1, np.int32)
    ts = rays["tmax"].copy()
    v0, v1, v2 = (scene.verts[scene.tris[:, k]].astype(np.float64) for k in range(3))
    e1, e2 = v1 - v0, v2 - v0
    for i, r in enumerate(rays):
        o, d = r["o"].astype(np.float64), r["d"].astype(np.float64)
        p = np.cross(d, e2)
        det = (e1 * p).sum(1)
        ok = np.abs(det) > 0
        inv = np.where(ok, 1.0 / np.where(ok, det, 1), 0)
        s = o - v0
        u = (s * p).sum(1) * inv
        q = np.cross(s, e1)
        v = (q * d).sum(1) * inv
        t = (q * e2).sum(1) * inv
        hit = ok & (u >= 0) & (v >= 0) & (u + v <= 1) & (t > r["tmin"]) & (t < ts[i])
        if hit.any():
            k = np.where(hit)[0][np.argmin(t[hit])]
            ids[i], ts[i] = k, t[k]
    return ids, ts


@pytest.mark.parametrize("kind", ["blob", "interior"])
def test_sbvh_builder_and_flatten(b2rt, oracle, kind):
    """Product builder output (Compact2 + Woop rows) is a valid BVH: the oracle tracers find the brute-force hits."""
    scene = b2rt.Scene.generate(kind, 1500, 9)
    bvh = b2rt.BVH(scene)
    arrays = bvh.arrays()
    st = bvh.stats
    assert st["numLeafNodes"] == st["numInnerNodes"] + 1 and st["numTris"] >= scene.getNumTriangles()
    assert arrays["nodes"].size == 16 * st["numInnerNodes"]
    assert arrays["woop"].size == 4 * (3 * st["numTris"] + st["numLeafNodes"])       # rows + one terminator per leaf
    # every triangle is referenced, Woop rows are the reference transform of that triangle
    woop = arrays["woop"].reshape(-1, 4)
    seen = set()
    i = 0
    while i < len(woop):
        if woop[i, 0].view(np.uint32) == 0x80000000:
            i += 1
            continue
        tri = int(arrays["triIndex"][i])
        seen.add(tri)
        if len(seen) % 97 == 0:
            v = scene.verts[scene.tris[tri]]
            assert oracle.woopify(v[0], v[1], v[2]).tobytes() == woop[i:i + 3].tobytes()
        i += 3
    assert len(seen) == scene.getNumTriangles()
    origin, m = b2rt.camera_matrices(scene.camera, 24, 16)
    i2p, _ = oracle.pixel_table(24, 16)
    rays, _, _ = oracle.raygen_primary(origin, m, 24, 16, scene.camera["far"], i2p)
    ids, ts = _brute_force(scene, rays)
    sc = {"verts": scene.verts, "tris": scene.tris}
    for got in (oracle.trace_mt(arrays, sc, rays), oracle.trace_woop(arrays, rays)):
        hit = ids >= 0
        assert ((got["id"] >= 0) == hit).mean() >= 0.995
        both = hit & (got["id"] >= 0)
        assert np.allclose(got["t"][both], ts[both], rtol=1e-4)


def test_bvhcache_file_round_trip(b2rt, oracle):
    """bvhcache format (Renderer.cc:176-211; CudaBVH.cc:113-116): S32 layout + 3 x (S64 size, bytes)."""
    scene = b2rt.Scene.generate("blob", 800, 3)
    bvh = b2rt.BVH(scene)
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "x.dat")
        bvh.serialize(p)
        ref = oracle.read_bvh(p)                       # independent reader of the reference format
        assert ref["layout"] == 5
        for k in ("nodes", "woop", "triIndex"):
            assert ref[k].tobytes() == np.asarray(bvh.arrays()[k]).tobytes()
        back = b2rt.BVH.load(p, scene)
        assert back.nodes.tobytes() == bvh.nodes.tobytes() and back.triIndex.tobytes() == bvh.triIndex.tobytes()
        # a file written by the reference's CudaBVH::serialize loads too
        g = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "room.npz")))
        oracle.write_bvh(p, {"layout": 5, "nodes": g["nodes"], "woop": g["woop"], "triIndex": g["triIndex"]})
        back = b2rt.BVH.load(p)
        assert back.woop.tobytes() == g["woop"].tobytes()
        with pytest.raises(b2rt.B2rtError):
            b2rt.BVH.load(os.path.join(d, "missing.dat"))


def test_raygen_batching(b2rt):
    """RayGen::batching (RayGen.cc:124-142)."""
    rg = b2rt.RayGen.__new__(b2rt.RayGen)
    rg.m_maxBatchSize, rg.m_aoStartIdx = 1 << 21, 0
    out, nb = [], True
    while True:
        cont, nb, lo, hi = rg.batching(786432, 32, nb)
        if not cont:
            break
        out.append((lo, hi))
    assert out[0] == (0, 65536) and out[-1][1] == 786432 and len(out) == 12
    assert all(a[1] == b[0] for a, b in zip(out, out[1:]))


def test_shard_range(b2rt):
    import importlib
    mg = importlib.import_module("gpu-ray-tracing_b200.multigpu")
    for n in (0, 1, 7, 64 * 1024 * 1024, 786431):
        for ws in (1, 2, 3, 8):
            parts = [mg.shard_range(n, r, ws) for r in range(ws)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1


def test_save_image_ppm(b2rt, tmp_path):
    """Row f1's headless image write: ABGR words -> binary PPM, bottom row first in memory (glDrawPixels convention)."""
    w, h = 5, 3
    px = np.zeros(w * h, dtype=np.uint32)
    px[0] = 0xFF0000FF           # red, bottom-left in the reference's window
    px[w * h - 1] = 0xFFFF0000   # blue, top-right
    px[w] = 0xFF00FF00           # green, second row from the bottom, left
    path = tmp_path / "frame.ppm"
    rgb = b2rt.save_image(px, w, h, str(path))
    raw = path.read_bytes()
    header = b"P6\n5 3\n255\n"
    assert raw.startswith(header) and len(raw) == len(header) + w * h * 3
    img = np.frombuffer(raw[len(header):], dtype=np.uint8).reshape(h, w, 3)
    assert (img == rgb).all()
    assert tuple(img[h - 1, 0]) == (255, 0, 0)        # bottom-left
    assert tuple(img[0, w - 1]) == (0, 0, 255)        # top-right
    assert tuple(img[h - 2, 0]) == (0, 255, 0)
    with pytest.raises(ValueError):
        b2rt.save_image(px[:-1], w, h, str(path))


@pytest.mark.parametrize("kind,target", [("blob", 3000), ("strands", 6000)])
def test_leaf_references_cover_every_triangle_row(b2rt, kind, target):
    """Input of the device Woop transform (b2rt_woopify): the (row, triangle) list recovered from the node array's leaf
    pointers names exactly the non-terminator rows of the host-built array, three rows apart inside a leaf."""
    scene = b2rt.Scene.generate(kind, target, 11)
    bvh = b2rt.BVH(scene)               # arrays() are views into the builder's memory: keep it alive
    a = bvh.arrays()
    rows, tris = b2rt.leaf_references(a["nodes"], a["triIndex"])
    woop = np.ascontiguousarray(a["woop"], np.float32).view(np.uint32).reshape(-1, 4)
    covered = np.zeros(len(woop), bool)
    for k in range(3):
        assert not covered[rows + k].any()
        covered[rows + k] = True
    term = ~covered
    assert (woop[term] == 0x80000000).all(), "rows outside the references are the leaf terminators"
    assert (woop[rows, 0] != 0x80000000).all()
    assert (tris == np.asarray(a["triIndex"])[rows]).all() and tris.min() >= 0 and tris.max() < len(scene.tris)
    assert len(rows) >= len(scene.tris)         # spatial splits only add references


def test_bvhcache_key_and_reference_written_file(b2rt, tmp_path):
    """Renderer::getCudaBVH names its cache file bvhcache/%08x.dat from hashes of the scene, the platform, the build parameters
    and the layout (Renderer.cc:176-186).  tests/golden/cache_keys.json holds the keys the reference's own code computed for
    the fixture scenes, tests/golden/bvhcache/<key>.dat a file the reference's CudaBVH::serialize wrote: the product computes
    the same names, reads that file, and writes the same bytes back."""
    import ctypes as C
    import json
    golden = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    keys = json.load(open(os.path.join(golden, "cache_keys.json")))
    lib = b2rt.lib()
    for name in ("room", "blob"):
        g = np.load(os.path.join(golden, name + ".npz"))
        v, t = np.ascontiguousarray(g["verts"], np.float32), np.ascontiguousarray(g["tris"], np.int32)
        key = C.c_uint32()
        assert lib.b2rt_bvhcache_key(v.ctypes.data_as(C.c_void_p), len(v), t.ctypes.data_as(C.c_void_p), len(t), None,
                                     C.c_float(1.0e-5), 1, 8, b2rt.LAYOUT_COMPACT2, C.byref(key)) == 0
        assert "%08x" % key.value == keys[name]["cache_key"]
        # any ingredient changes the name
        other = C.c_uint32()
        lib.b2rt_bvhcache_key(v.ctypes.data_as(C.c_void_p), len(v), t.ctypes.data_as(C.c_void_p), len(t), None, C.c_float(1.0e-4), 1, 8,
                              b2rt.LAYOUT_COMPACT2, C.byref(other))
        assert other.value != key.value
    buf = C.create_string_buffer(256)
    assert lib.b2rt_bvhcache_path(os.path.join(golden, "bvhcache").encode(), int(keys["room"]["cache_key"], 16), buf, 256) == 0
    path = buf.value.decode()
    assert os.path.exists(path), "the reference-written cache file is not where the product looks for it"
    bvh = b2rt.BVH.load(path)
    g = np.load(os.path.join(golden, "room.npz"))
    assert bvh.nodes.tobytes() == g["nodes"].tobytes() and bvh.woop.tobytes() == g["woop"].tobytes()
    assert bvh.triIndex.tobytes() == g["triIndex"].tobytes()
    out = str(tmp_path / "again.dat")
    bvh.serialize(out)
    assert open(out, "rb").read() == open(path, "rb").read()
    assert lib.b2rt_bvhcache_path(b"x" * 300, 1, buf, 256) != 0          # a path that does not fit is refused
