"""The private node array b2rt_set_bvh derives from the caller's Compact2 nodes (api.cu: relayoutNodes + packLeafCounts),
checked on the CPU through the host-only entry b2rt_private_nodes: the tree is the caller's tree (same boxes, same
leaves), every leaf reference carries its triangle count, and a scalar traversal of the private form that trusts the
counts (never looks at a terminator row) returns the oracle's hits on the caller's arrays."""
import ctypes as C
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ROW_BITS = 26
SENTINEL = 0x76543210


def private_nodes(b2rt, nodes, woop):
    lib = b2rt.lib()
    nodes = np.ascontiguousarray(nodes, np.float32).reshape(-1)
    woop = np.ascontiguousarray(woop, np.float32).reshape(-1)
    nbytes, top, counted = C.c_size_t(0), C.c_int(0), C.c_int(0)
    args = (nodes.ctypes.data_as(C.c_void_p), nodes.nbytes, woop.ctypes.data_as(C.c_void_p), woop.nbytes)
    assert lib.b2rt_private_nodes(*args, None, 0, C.byref(nbytes), C.byref(top), C.byref(counted)) == 0
    out = np.zeros(nbytes.value // 4, np.float32)
    assert lib.b2rt_private_nodes(*args, out.ctypes.data_as(C.c_void_p), out.nbytes, C.byref(nbytes), C.byref(top), C.byref(counted)) == 0
    return out, top.value, counted.value


def leaves_of(nodes_i32, decode):
    """[(row, count)] of every leaf reachable from the root, in depth-first order."""
    out, stack = [], [0]
    while stack:
        a = stack.pop()
        if a < 0:
            out.append(decode(a))
            continue
        n = nodes_i32[a // 4]
        stack.append(int(n[13])); stack.append(int(n[12]))
    return out


def count_to_terminator(woop_bits, row):
    n = 0
    while woop_bits[row] != 0x80000000:
        n += 1
        row += 3
    return n


def trace_counted(priv, woop, ray):
    """Closest hit by ordered stack traversal of the PRIVATE form; leaves run `count` triangles, no terminator test."""
    o, d, tmin, hit_t = ray["o"].astype(np.float32), ray["d"].astype(np.float32), np.float32(ray["tmin"]), np.float32(ray["tmax"])
    eps = np.float32(2.0 ** -80)
    dd = np.where(np.abs(d) > eps, d, np.copysign(eps, d)).astype(np.float32)
    idir = (np.float32(1.0) / dd).astype(np.float32)
    ood = (o * idir).astype(np.float32)
    nodes = priv.reshape(-1, 16)
    ni = priv.view(np.int32).reshape(-1, 16)
    w = woop.reshape(-1, 4)
    hit, stack, addr = -1, [SENTINEL], 0
    while addr != SENTINEL:
        if addr >= 0:
            n = nodes[addr // 4]
            ch = ni[addr // 4]
            res = []
            for c in range(2):
                lo = np.array([n[4 * c + 0], n[4 * c + 2], n[8 + 2 * c]], np.float32)
                hi = np.array([n[4 * c + 1], n[4 * c + 3], n[9 + 2 * c]], np.float32)
                t0 = lo * idir - ood
                t1 = hi * idir - ood
                cmin = max(np.minimum(t0, t1).max(), tmin)
                cmax = min(np.maximum(t0, t1).min(), hit_t)
                res.append((cmax >= cmin, cmin))
            (h0, m0), (h1, m1) = res
            if not h0 and not h1:
                addr = stack.pop()
            else:
                addr = int(ch[12]) if h0 else int(ch[13])
                if h0 and h1:
                    far = int(ch[13])
                    if m1 < m0:
                        far, addr = addr, int(ch[13])
                    stack.append(far)
        else:
            ref = ~addr
            row, cnt = ref & ((1 << ROW_BITS) - 1), ref >> ROW_BITS
            for k in range(cnt):
                r = row + 3 * k
                v0, v1, v2 = w[r], w[r + 1], w[r + 2]
                oz = v0[3] - np.dot(o, v0[:3])
                t = np.float32(oz / np.dot(d, v0[:3]))
                if tmin < t < hit_t:
                    u = np.float32(v1[3] + np.dot(o, v1[:3]) + t * np.dot(d, v1[:3]))
                    if u >= 0:
                        v = np.float32(v2[3] + np.dot(o, v2[:3]) + t * np.dot(d, v2[:3]))
                        if v >= 0 and u + v <= 1:
                            hit_t, hit = t, r
            addr = stack.pop()
    return hit, hit_t


@pytest.mark.parametrize("case", ["blob", "room"])
def test_private_form_of_reference_built_bvh(b2rt, oracle, case):
    g = dict(np.load(os.path.join(GOLDEN, case + ".npz")))
    nodes, woop, tri = g["nodes"], g["woop"], g["triIndex"]
    priv, top, counted = private_nodes(b2rt, nodes, woop)
    assert counted == 1 and 0 < top <= 4096
    assert priv.size == np.asarray(nodes).size                   # no leaf of more than 31 triangles here: nothing appended
    bits = np.ascontiguousarray(woop, np.float32).reshape(-1, 4)[:, 0].view(np.uint32)
    orig = leaves_of(np.ascontiguousarray(nodes, np.float32).view(np.int32).reshape(-1, 16), lambda a: (~a, None))
    mine = leaves_of(priv.view(np.int32).reshape(-1, 16), lambda a: ((~a) & ((1 << ROW_BITS) - 1), (~a) >> ROW_BITS))
    # same leaves in the same depth-first order, each with the count the terminator row implies
    assert [r for r, _ in orig] == [r for r, _ in mine]
    for row, cnt in mine:
        assert 1 <= cnt <= 31 and cnt == count_to_terminator(bits, row)
    # boxes travel with their nodes: multiset of the 12 box floats per node is unchanged
    a = np.sort(np.ascontiguousarray(nodes, np.float32).reshape(-1, 16)[:, :12].view(np.uint32), axis=0)
    b = np.sort(priv.reshape(-1, 16)[:, :12].view(np.uint32), axis=0)
    assert (a == b).all()
    # counted traversal of the private form == the oracle's traversal of the caller's arrays
    rays = np.concatenate([g["prim_rays"][::97], g["diffuse_rays"][::193]])[:160]
    ref = oracle.trace_woop({"nodes": nodes, "woop": woop, "triIndex": tri}, rays)
    triIndex = np.asarray(tri).reshape(-1)
    agree = 0
    for i, r in enumerate(rays):
        hit, t = trace_counted(priv, np.ascontiguousarray(woop, np.float32), r)
        rid = -1 if hit < 0 else int(triIndex[hit])
        if rid == ref["id"][i]:
            agree += 1
            if rid >= 0:
                assert abs(t - ref["t"][i]) <= 1e-5 * abs(ref["t"][i])
    assert agree >= len(rays) - 1


def test_big_leaf_is_cut_into_pieces(b2rt):
    """One inner node over a 70-triangle leaf and a 2-triangle leaf: the big leaf becomes a chain of 31 + 31 + 8."""
    n_big, n_small = 70, 2
    rows = 3 * n_big + 1 + 3 * n_small + 1
    woop = np.ones((rows, 4), np.float32)
    woop[3 * n_big] = 0
    woop[3 * n_big, 0] = np.uint32(0x80000000).view(np.float32)
    woop[rows - 1] = 0
    woop[rows - 1, 0] = np.uint32(0x80000000).view(np.float32)
    node = np.zeros(16, np.float32)
    node[0:4] = [-1, 1, -2, 2]          # c0 lox hix loy hiy
    node[4:8] = [3, 4, 5, 6]
    node[8:12] = [-3, 3, 7, 8]          # c0 loz hiz, c1 loz hiz
    node.view(np.int32)[12:16] = [~0, ~(3 * n_big + 1), 0, 0]
    priv, top, counted = private_nodes(b2rt, node, woop)
    assert counted == 1 and priv.size == 3 * 16
    pi = priv.view(np.int32).reshape(-1, 16)
    pf = priv.reshape(-1, 16)
    dec = lambda a: ((~a) & ((1 << ROW_BITS) - 1), (~a) >> ROW_BITS)
    assert dec(int(pi[0, 13])) == (3 * n_big + 1, n_small)
    assert pi[0, 12] == 4                                        # link to the first appended node
    assert dec(int(pi[1, 12])) == (0, 31) and pi[1, 13] == 8
    assert dec(int(pi[2, 12])) == (93, 31) and dec(int(pi[2, 13])) == (186, 8)
    for k in (1, 2):                                             # both children repeat the big leaf's box
        assert list(pf[k, 0:4]) == [-1, 1, -2, 2] and list(pf[k, 4:8]) == [-1, 1, -2, 2]
        assert list(pf[k, 8:12]) == [-3, 3, -3, 3]
    got = leaves_of(pi, dec)
    assert got == [(0, 31), (93, 31), (186, 8), (3 * n_big + 1, n_small)]
