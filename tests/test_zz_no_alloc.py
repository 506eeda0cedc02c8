"""b2rt_count_hits_static (include/b2rt.h, csrc/static_scratch.cu): the hit count for callers that cannot allocate -- scratch in a
file-scope device array, result in a device word the caller passes.  No counterpart in the reference.
Called through the C ABI and compared with the existing entry point b2rt_count_hits: integer work, so the
tolerance is the one the other count tests use -- exact equality.  Two calls in a row on the same stream share the scratch;
the second batch is SMALLER than the first, so partial counts the first call left behind would show.

Kept in a file of its own that pytest collects last.  Passes alone in a fresh process and
after the other tests: it needs nothing from them but the built library."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _records(torch, n, seed, hit_fraction):
    """n RayResult records (16 B: id, t, padA, padB) on the device: id >= 0 for about hit_fraction of them, -1 otherwise."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    rec = torch.zeros((max(n, 1), 4), dtype=torch.int32, device="cuda")
    ids = torch.randint(0, 1 << 20, (max(n, 1),), generator=g, device="cuda", dtype=torch.int32)
    miss = torch.rand((max(n, 1),), generator=g, device="cuda") >= hit_fraction
    ids[miss] = -1
    rec[:, 0] = ids
    rec[:, 1] = 0x3F800000
    return rec[:n] if n else rec[:0]


def _static(b2rt, torch, rec, n, out):
    rc = b2rt.lib().b2rt_count_hits_static(C.c_void_p(rec.data_ptr()), n, C.c_void_p(out.data_ptr()),
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, b2rt.lib().b2rt_last_error(None)


def test_count_hits_static_matches_count_hits_on_two_calls_in_a_row(b2rt, ctx):
    import torch
    torch.cuda.set_device(0)
    out = torch.full((2,), -7, dtype=torch.int64, device="cuda")            # the caller's output words, allocated up front
    first = _records(torch, 3 * 1024 * 1024 + 1234, 11, 0.6)                # 769 blocks, ragged last tile
    second = _records(torch, 70001, 12, 0.25)                               # 18 blocks: far fewer partial counts than the first call wrote
    torch.cuda.synchronize()
    # two calls back to back on one stream, no synchronisation and no allocation in between
    _static(b2rt, torch, first, first.shape[0], out[0:1])
    _static(b2rt, torch, second, second.shape[0], out[1:2])
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    want = [b2rt.countHits(first, ctx), b2rt.countHits(second, ctx)]        # the existing entry point, b2rt_count_hits
    assert got.tolist() == want
    assert want == [int((first[:, 0] >= 0).sum().item()), int((second[:, 0] >= 0).sum().item())]
    # and again in the other order: the scratch keeps the last call's contents, the answers do not change
    _static(b2rt, torch, second, second.shape[0], out[0:1])
    _static(b2rt, torch, first, first.shape[0], out[1:2])
    torch.cuda.synchronize()
    assert out.cpu().numpy().tolist() == want[::-1]


def test_count_hits_static_edges_and_limit(b2rt, ctx):
    import torch
    torch.cuda.set_device(0)
    lib = b2rt.lib()
    out = torch.full((1,), -7, dtype=torch.int64, device="cuda")
    rec = _records(torch, 4096 * 3, 5, 1.0)
    # an empty batch counts zero hits
    _static(b2rt, torch, rec, 0, out)
    torch.cuda.synchronize()
    assert int(out.item()) == 0
    # exactly whole tiles, every record a hit; one record; all misses
    _static(b2rt, torch, rec, rec.shape[0], out)
    torch.cuda.synchronize()
    assert int(out.item()) == rec.shape[0] == b2rt.countHits(rec, ctx)
    _static(b2rt, torch, rec, 1, out)
    torch.cuda.synchronize()
    assert int(out.item()) == 1
    none = _records(torch, 5000, 6, 0.0)
    _static(b2rt, torch, none, none.shape[0], out)
    torch.cuda.synchronize()
    assert int(out.item()) == 0 == b2rt.countHits(none, ctx)
    # beyond the stated maximum: refused with a clear message, nothing launched, the output untouched
    out.fill_(-7)
    torch.cuda.synchronize()
    too_many = b2rt._lib.COUNT_HITS_STATIC_MAX + 1
    rc = lib.b2rt_count_hits_static(C.c_void_p(rec.data_ptr()), too_many, C.c_void_p(out.data_ptr()), None)
    assert rc == -1
    msg = lib.b2rt_last_error(None).decode()
    assert "exceed" in msg and str(b2rt._lib.COUNT_HITS_STATIC_MAX) in msg
    torch.cuda.synchronize()
    assert int(out.item()) == -7
    # bad arguments
    assert lib.b2rt_count_hits_static(None, 16, C.c_void_p(out.data_ptr()), None) == -1
    assert lib.b2rt_count_hits_static(C.c_void_p(rec.data_ptr()), 16, None, None) == -1
    assert lib.b2rt_count_hits_static(C.c_void_p(rec.data_ptr()), -1, C.c_void_p(out.data_ptr()), None) == -1


def test_count_hits_static_at_the_maximum(b2rt, ctx):
    import torch
    torch.cuda.set_device(0)
    n = b2rt._lib.COUNT_HITS_STATIC_MAX                                      # 16 Mi records = 256 MB: every partial-count slot is used
    rec = _records(torch, n, 21, 0.5)
    out = torch.zeros((1,), dtype=torch.int64, device="cuda")
    _static(b2rt, torch, rec, n, out)
    torch.cuda.synchronize()
    assert int(out.item()) == b2rt.countHits(rec, ctx) == int((rec[:, 0] >= 0).sum().item())
