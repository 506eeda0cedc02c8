"""Scalar Python statement of the 8-wide compressed traversal (test infrastructure): the same node
decode, octant ordering and hit-mask logic as csrc/trace_wide.cu, one ray at a time, float64 boxes.
Used by tests/test_wide_bvh.py to check the host collapse (csrc/host/wide_bvh.cpp) without a GPU."""
import ctypes as C

import numpy as np


def build(b2rt, arrays):
    """b2rt_wide_build -> dict(nodes uint32 [n,20], woop float32 [t,3,4], triIndex int32 [t], stats)."""
    lib = b2rt.lib()
    h = C.c_void_p()
    n = np.ascontiguousarray(arrays["nodes"], np.float32)
    w = np.ascontiguousarray(arrays["woop"], np.float32)
    t = np.ascontiguousarray(arrays["triIndex"], np.int32)
    rc = lib.b2rt_wide_build(n.ctypes.data_as(C.c_void_p), n.nbytes, w.ctypes.data_as(C.c_void_p), w.nbytes,
                             t.ctypes.data_as(C.c_void_p), t.nbytes, C.byref(h))
    assert rc == 0, "b2rt_wide_build failed (%d)" % rc
    try:
        pn, pw, pt = C.c_void_p(), C.c_void_p(), C.c_void_p()
        bn, bw, bt = C.c_size_t(), C.c_size_t(), C.c_size_t()
        stats = (C.c_int * 3)()
        rc = lib.b2rt_wide_arrays(h, C.byref(pn), C.byref(bn), C.byref(pw), C.byref(bw), C.byref(pt), C.byref(bt), stats)
        assert rc == 0
        nodes = np.ctypeslib.as_array(C.cast(pn, C.POINTER(C.c_uint32)), (bn.value // 4,)).copy().reshape(-1, 20)
        woop = np.ctypeslib.as_array(C.cast(pw, C.POINTER(C.c_float)), (bw.value // 4,)).copy()
        tri = np.ctypeslib.as_array(C.cast(pt, C.POINTER(C.c_int32)), (bt.value // 4,)).copy()
    finally:
        lib.b2rt_wide_free(h)
    ntri = len(tri)
    return {"nodes": nodes, "woop": woop[:ntri * 12].reshape(ntri, 3, 4), "triIndex": tri,
            "stats": {"nodes": stats[0], "tris": stats[1], "depth": stats[2]}}


def _node(nodes, i):
    w = nodes[i]
    p = w[0:3].view(np.float32).astype(np.float64)
    eb = [(int(w[3]) >> (8 * k)) & 0xFF for k in range(3)]
    imask = int(w[3]) >> 24
    meta = np.frombuffer(w[6:8].tobytes(), np.uint8)
    q = np.frombuffer(w[8:20].tobytes(), np.uint8).reshape(6, 8).astype(np.float64)      # qlo xyz, qhi xyz
    step = np.array([2.0 ** (e - 142) for e in eb])
    lo = p[:, None] + step[:, None] * q[0:3]
    hi = p[:, None] + step[:, None] * q[3:6]
    return imask, int(w[4]), int(w[5]), meta, lo, hi


def _fma(a, b, c):
    """float32 fused multiply-add (the product of two float32 is exact in float64)."""
    return np.float32(np.float64(a) * np.float64(b) + np.float64(c))


def trace(wide, ray, any_hit=False):
    """Closest (or any) hit of one ray: returns (private triangle index or -1, t, nodes visited, triangles tested)."""
    nodes, woop = wide["nodes"], wide["woop"]
    o = np.array(ray["o"], np.float64)
    d = np.array(ray["d"], np.float64)
    o32, d32 = np.array(ray["o"], np.float32), np.array(ray["d"], np.float32)
    tmin, hit_t, hit = float(ray["tmin"]), float(ray["tmax"]), -1
    eps = 2.0 ** -80
    dd = np.where(np.abs(d) > eps, d, np.copysign(eps, d))
    idir = 1.0 / dd
    oct_inv = (0 if d[0] < 0 else 4) | (0 if d[1] < 0 else 2) | (0 if d[2] < 0 else 1)
    stack = []
    G = (0, 0x80000000)
    n_nodes = n_tris = 0
    while True:
        T = (0, 0)
        if G[1] > 0x00FFFFFF:
            bit = G[1].bit_length() - 1
            G = (G[0], G[1] & ~(1 << bit))
            if G[1] > 0x00FFFFFF:
                stack.append(G)
            slot = (bit - 24) ^ oct_inv
            rel = bin(G[1] & 0xFF & ((1 << slot) - 1)).count("1")
            imask, child_base, tri_base, meta, lo, hi = _node(nodes, G[0] + rel)
            n_nodes += 1
            hitmask = 0
            for s in range(8):
                m = int(meta[s])
                if m == 0:
                    continue
                t0 = (lo[:, s] - o) * idir
                t1 = (hi[:, s] - o) * idir
                near = np.minimum(t0, t1)
                far = np.maximum(t0, t1)
                cmin = max(near.max(), tmin)
                cmax = min(far.min(), hit_t)
                if cmin <= cmax:
                    inner = (m & 0x18) == 0x18
                    bit_index = ((m ^ oct_inv) & 31) if inner else (m & 31)
                    hitmask |= (m >> 5) << bit_index
            G = (child_base, (hitmask & 0xFF000000) | imask)
            T = (tri_base, hitmask & 0x00FFFFFF)
        else:
            T, G = G, (0, 0)
        tm = T[1]
        while tm:
            k = tm.bit_length() - 1
            tm &= ~(1 << k)
            tri = T[0] + k
            n_tris += 1
            # the Woop test in float32 with the kernel's / oracle's FMA order (kepler_dynamic_fetch.cu:336-379)
            r0, r1, r2 = woop[tri]
            oz = _fma(-o32[2], r0[2], _fma(-o32[1], r0[1], _fma(-o32[0], r0[0], r0[3])))
            dz = _fma(d32[2], r0[2], _fma(d32[1], r0[1], np.float32(d32[0] * r0[0])))
            with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
                t = np.float32(oz * (np.float32(1.0) / dz))
            if tmin < t < hit_t:
                u = _fma(t, _fma(d32[2], r1[2], _fma(d32[1], r1[1], np.float32(d32[0] * r1[0]))),
                         _fma(o32[2], r1[2], _fma(o32[1], r1[1], _fma(o32[0], r1[0], r1[3]))))
                if u >= 0:
                    v = _fma(t, _fma(d32[2], r2[2], _fma(d32[1], r2[1], np.float32(d32[0] * r2[0]))),
                             _fma(o32[2], r2[2], _fma(o32[1], r2[1], _fma(o32[0], r2[0], r2[3]))))
                    if v >= 0 and np.float32(u + v) <= 1:
                        hit_t, hit = float(t), tri
                        if any_hit:
                            return hit, hit_t, n_nodes, n_tris
        if G[1] <= 0x00FFFFFF:
            if not stack:
                break
            G = stack.pop()
    return hit, hit_t, n_nodes, n_tris
