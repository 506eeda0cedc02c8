"""B2RT_TRACE_WATERTIGHT (include/b2rt.h; common.cuh:woopAccept): edge-tolerant Woop test + order-independent closest hit.

The reference's test is exact (u >= 0, v >= 0, u + v <= 1; kepler_dynamic_fetch.cu:349-364, Util.cc:99-127), so a ray aimed
at the edge two triangles share can fail both by one rounding and pass through a closed surface.  With the flag:
  * no leaks: rays aimed EXACTLY at shared edges and vertices -- a quad split along its diagonal, a fan around one vertex, and
    every vertex and edge midpoint of a closed surface seen from inside -- all hit (wherever Intersect::RayTriangle, the
    reference's CPU test, or plain geometry says they must);
  * the same record for a ray wherever it sits in the batch: ids and t bit-identical under a permutation of the ray buffer
    (without the flag ids flip between the two triangles of an exact tie, tests/test_gpu_fullsize.py tolerates 0.05 %);
  * away from edges nothing changes: ids equal the exact test's.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _trace(b2rt, oracle, ctx, scene, rays, watertight, variant=-1, closest=True):
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    tracer.watertight = watertight
    rb = b2rt.RayBuffer(closestHit=closest)
    rb.setRays(rays)
    rb.setNeedClosestHit(closest)
    ctx.set_trace_config(variant=variant)
    try:
        tracer.traceBatch(rb)
    finally:
        ctx.set_trace_config()
    return rb.getResults().copy()


def _rays_at(oracle, targets, origins):
    """Unnormalised rays from origins[i] THROUGH targets[i] (t = 1 at the target), interval (0, 4)."""
    rays = np.zeros(len(targets), oracle.RAY_DTYPE)
    rays["o"] = origins
    rays["d"] = (targets - origins).astype(np.float32)
    rays["tmin"] = 0.0
    rays["tmax"] = 4.0
    return rays


def _closed_bumpy_sphere(b2rt, rings, segs):
    """UV sphere with two pole vertices and a welded seam, radius 1 + low-frequency bumps: 2 * segs * (rings - 1) triangles."""
    th = np.pi * np.arange(1, rings) / rings
    ph = 2 * np.pi * np.arange(segs) / segs
    T, P = np.meshgrid(th, ph, indexing="ij")
    r = 1.0 + 0.12 * np.sin(3 * T + 0.4) * np.cos(4 * P + 0.9) + 0.05 * np.cos(5 * T) * np.sin(2 * P)
    body = np.stack([r * np.sin(T) * np.cos(P), r * np.sin(T) * np.sin(P), r * np.cos(T)], -1).reshape(-1, 3)
    verts = np.concatenate([[[0, 0, 1.07]], body, [[0, 0, -0.95]]]).astype(np.float32)
    idx = lambda i, j: 1 + i * segs + (j % segs)
    tris = []
    for j in range(segs):
        tris.append([0, idx(0, j), idx(0, j + 1)])
        tris.append([len(verts) - 1, idx(rings - 2, j + 1), idx(rings - 2, j)])
    i, j = np.meshgrid(np.arange(rings - 2), np.arange(segs), indexing="ij")
    a, b, c, d = idx(i, j), idx(i + 1, j), idx(i + 1, j + 1), idx(i, j + 1)
    tris = np.concatenate([np.array(tris), np.stack([a, b, c], -1).reshape(-1, 3), np.stack([a, c, d], -1).reshape(-1, 3)]).astype(np.int32)
    e = np.sort(np.concatenate([tris[:, [0, 1]], tris[:, [1, 2]], tris[:, [2, 0]]]), axis=1)
    _, cnt = np.unique(e, axis=0, return_counts=True)
    assert (cnt == 2).all(), "the test surface is not closed"
    return b2rt.Scene(verts, tris)


def _directions(n, seed):
    rng = np.random.default_rng(seed)
    v = rng.normal(size=(n, 3))
    v /= np.linalg.norm(v, axis=1)[:, None]
    return v


@pytest.mark.parametrize("variant", [-1, 1, 108])
def test_quad_diagonal_and_fan_do_not_leak(b2rt, oracle, ctx, variant):
    rng = np.random.default_rng(3)
    # a quad in a tilted plane, split along its diagonal v0-v2; awkward coordinates so that nothing is exactly representable
    base = np.array([[0.1371, -0.731, 0.2113], [1.3173, -0.6531, 0.4177], [1.2291, 0.5713, 0.7919], [0.0489, 0.4934, 0.5855]], np.float32)
    quad = b2rt.Scene(base, np.array([[0, 1, 2], [0, 2, 3]], np.int32))
    n = 20000
    s = rng.random(n).astype(np.float32)
    on_diag = (base[0] * (1 - s)[:, None] + base[2] * s[:, None]).astype(np.float32)            # points ON the shared edge
    normal = np.cross(base[1] - base[0], base[3] - base[0])
    normal /= np.linalg.norm(normal)
    dirs = _directions(n, 4)
    dirs[(dirs @ normal) < 0] *= -1                                                              # all from one side
    dirs = dirs[np.abs(dirs @ normal) > 0.05][: n]
    on_diag = on_diag[: len(dirs)]
    rays = _rays_at(oracle, on_diag, (on_diag + 1.7 * dirs).astype(np.float32))
    exact = _trace(b2rt, oracle, ctx, quad, rays, False, variant)
    tight = _trace(b2rt, oracle, ctx, quad, rays, True, variant)
    assert (tight["id"] >= 0).all(), "%d rays aimed at the shared diagonal leak through the quad" % int((tight["id"] < 0).sum())
    assert np.abs(tight["t"] - 1.0).max() < 1e-4
    print("quad diagonal, variant %d: exact test leaks %d of %d rays, watertight 0" % (variant, int((exact["id"] < 0).sum()), len(rays)))

    # a fan of 11 triangles around one vertex, nearly flat; rays aimed at the hub and along the spokes
    k = 11
    ang = np.linspace(0, 2 * np.pi, k, endpoint=False) + 0.123
    # (ripple of 0.04 over a radius of 1.2: no spoke is ever a silhouette edge for rays steeper than 17 degrees, so a ray through a
    #  point of the fan cannot geometrically pass it by)
    rim = np.stack([1.3 * np.cos(ang) + 0.317, 1.1 * np.sin(ang) - 0.211, 0.04 * np.sin(3 * ang) + 0.05], 1)
    hub = np.array([[0.3311, -0.1973, 0.0927]])
    verts = np.concatenate([hub, rim]).astype(np.float32)
    tris = np.array([[0, 1 + i, 1 + (i + 1) % k] for i in range(k)], np.int32)
    fan = b2rt.Scene(verts, tris)
    m = 20000
    spoke = rng.integers(0, k, m)
    s = (rng.random(m) * 0.98).astype(np.float32)
    s[: m // 4] = 0.0                                                                            # a quarter of them: the hub itself
    targets = (verts[0] * (1 - s)[:, None] + verts[1 + spoke] * s[:, None]).astype(np.float32)
    dirs = _directions(m, 5)
    dirs[dirs[:, 2] < 0] *= -1
    dirs = dirs[dirs[:, 2] > 0.5]                                                                 # steep: no spoke is a silhouette edge
    targets = targets[: len(dirs)]
    rays = _rays_at(oracle, targets, (targets + 1.9 * dirs).astype(np.float32))
    exact = _trace(b2rt, oracle, ctx, fan, rays, False, variant)
    tight = _trace(b2rt, oracle, ctx, fan, rays, True, variant)
    assert (tight["id"] >= 0).all(), "%d rays aimed at the fan's hub / spokes leak" % int((tight["id"] < 0).sum())
    print("fan hub and spokes, variant %d: exact test leaks %d of %d rays, watertight 0" % (variant, int((exact["id"] < 0).sum()), len(rays)))


@pytest.mark.parametrize("variant", [-1, 108])
def test_closed_surface_from_inside_has_no_holes(b2rt, oracle, ctx, variant):
    """A closed, star-shaped genus-0 surface (every edge has exactly two triangles: asserted): from a point inside, a ray through
    ANY of its vertices or edge midpoints must hit something -- with the exact test some slip between the triangles."""
    scene = _closed_bumpy_sphere(b2rt, 96, 192)
    verts, tris = scene.verts, scene.tris
    inside = np.array([0.013, -0.021, 0.017], np.float32)
    edges = np.concatenate([tris[:, [0, 1]], tris[:, [1, 2]], tris[:, [2, 0]]])
    mid = ((verts[edges[:, 0]].astype(np.float64) + verts[edges[:, 1]]) * 0.5).astype(np.float32)
    targets = np.concatenate([verts, mid])
    origins = np.broadcast_to(inside, targets.shape).copy()
    rays = _rays_at(oracle, targets, origins)
    rays["tmax"] = 8.0
    exact = _trace(b2rt, oracle, ctx, scene, rays, False, variant)
    tight = _trace(b2rt, oracle, ctx, scene, rays, True, variant)
    assert (tight["id"] >= 0).all(), "%d of %d rays escape the closed surface" % (int((tight["id"] < 0).sum()), len(rays))
    print("closed surface, variant %d: exact test lets %d of %d rays escape, watertight 0" % (variant, int((exact["id"] < 0).sum()), len(rays)))
    # any hit sees the same thing
    anyhit = _trace(b2rt, oracle, ctx, scene, rays, True, variant, closest=False)
    assert (anyhit["id"] >= 0).all()


@pytest.mark.parametrize("variant", [-1, 1, 108])
def test_watertight_results_do_not_depend_on_the_order_of_the_rays(b2rt, oracle, ctx, variant):
    """Interior scene (walls, furniture: doubled coplanar faces, thousands of exactly tied hits): primary + diffuse rays, traced
    in buffer order and shuffled.  With the flag the records are bit-identical; the exact test keeps whichever tied triangle a
    lane meets first."""
    scene = b2rt.Scene.generate("interior", 30000, 7)
    bvh = b2rt.BVH(scene)
    origin, m = b2rt.camera_matrices(scene.camera, 320, 240)
    i2p, _ = oracle.pixel_table(320, 240)
    prim, _, _ = oracle.raygen_primary(origin, m, 320, 240, scene.camera["far"], i2p)
    res = oracle.trace_woop(bvh.arrays(), prim)
    sec, _, _ = oracle.raygen_ao(0, len(prim), 2, scene.camera["far"], 4242, prim, res, scene.normals)
    rays = np.concatenate([prim, sec[sec["tmax"] >= 0]])
    perm = np.random.default_rng(9).permutation(len(rays))
    flips = {}
    for wt in (False, True):
        a = _trace(b2rt, oracle, ctx, scene, rays, wt, variant)
        b = _trace(b2rt, oracle, ctx, scene, rays[perm], wt, variant)
        back = np.empty_like(b)
        back[perm] = b
        flips[wt] = int((a["id"] != back["id"]).sum())
        if wt:
            assert flips[wt] == 0, "%d ids change when the ray buffer is shuffled" % flips[wt]
            assert (a["t"] == back["t"]).all() and (a["padA"] == back["padA"]).all() and (a["padB"] == back["padB"]).all()
            tight = a
        else:
            exact = a
    print("permutation, variant %d: %d of %d ids flip with the exact test, 0 with the watertight one" % (variant, flips[False], len(rays)))
    # away from ties the flag changes nothing: same hit-or-miss everywhere it matters, same t where the ids agree
    same = tight["id"] == exact["id"]
    assert same.mean() >= 0.95                                   # the rest are the scene's exact ties: lower id wins instead of first met
    assert (tight["t"][same] == exact["t"][same]).all()
    both = (tight["id"] >= 0) & (exact["id"] >= 0)
    assert (tight["t"][both] <= exact["t"][both] * (1 + 2e-6)).all()          # the tolerance can add a NEARER hit (an edge-on triangle), never lose one
    assert ((exact["id"] >= 0) <= (tight["id"] >= 0)).all()      # the tolerance only ever ADDS hits
    # against BVH::trace's own test (restatement b2o_trace_mt): agreement with and without the flag
    mt = oracle.trace_mt({k: np.ascontiguousarray(v) for k, v in bvh.arrays().items() if k != "layout"},
                         {"verts": scene.verts, "tris": scene.tris}, rays)
    print("   hit / miss agreement with BVH::trace: exact %.6f, watertight %.6f; id agreement: exact %.6f, watertight %.6f" % (
        ((exact["id"] >= 0) == (mt["id"] >= 0)).mean(), ((tight["id"] >= 0) == (mt["id"] >= 0)).mean(),
        (exact["id"] == mt["id"]).mean(), (tight["id"] == mt["id"]).mean()))
    assert ((tight["id"] >= 0) == (mt["id"] >= 0)).mean() >= 0.9999
