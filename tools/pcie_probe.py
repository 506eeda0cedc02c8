"""Pinned host<->device copy rates of the box (what bounds b2rt_trace_host): one direction, then both at once."""
import torch, time
dev = torch.device("cuda:0")
for mb in (8, 25, 64, 256):
    n = mb << 20
    h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
    d_a = torch.empty(n, dtype=torch.uint8, device=dev)
    d_b = torch.empty(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    def timed(fn, reps=10):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps): fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps
    def h2d():
        with torch.cuda.stream(s1): d_a.copy_(h_in, non_blocking=True)
    def d2h():
        with torch.cuda.stream(s2): h_out.copy_(d_b, non_blocking=True)
    def both():
        h2d(); d2h()
    t1, t2, t3 = timed(h2d), timed(d2h), timed(both)
    print("%4d MiB: H2D %.1f GB/s  D2H %.1f GB/s  both at once %.1f + %.1f GB/s" % (mb, n / t1 / 1e9, n / t2 / 1e9, n / t3 / 1e9, n / t3 / 1e9))
