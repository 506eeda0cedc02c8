"""Pinned host<->device copy rates with 1..N GPUs copying AT THE SAME TIME (what bounds b2rt_trace_host at N ranks): one
process per GPU, started together; with --bind each process first pins itself to its share of the GPU-local cores
(multigpu.bind_to_local_cores) so that its pinned buffer is first-touched there.
usage: python tools/pcie_probe_all.py N [--bind]      (prints one line per GPU and the aggregate)"""
import importlib, os, subprocess, sys, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

def worker(rank, n, bind, start_at):
    sys.path.insert(0, ROOT)
    if bind:
        mg = importlib.import_module("gpu-ray-tracing_b200.multigpu")
        mg.bind_to_local_cores(rank, n)
    import torch
    torch.cuda.set_device(rank)
    nbytes = 256 << 20
    h_in = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_a = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    d_b = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    def both():
        with torch.cuda.stream(s1): d_a.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s2): h_out.copy_(d_b, non_blocking=True)
    both(); torch.cuda.synchronize()
    while time.time() < start_at: pass
    reps = 20
    t0 = time.perf_counter()
    for _ in range(reps): both()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    print("gpu %d: H2D + D2H at once %.1f + %.1f GB/s" % (rank, nbytes / dt / 1e9, nbytes / dt / 1e9), flush=True)

if __name__ == "__main__":
    if sys.argv[1] == "--worker":
        worker(int(sys.argv[2]), int(sys.argv[3]), sys.argv[4] == "1", float(sys.argv[5]))
    else:
        n = int(sys.argv[1]); bind = "--bind" in sys.argv
        start = time.time() + 25.0
        ps = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", str(r), str(n), "1" if bind else "0", str(start)],
                               stdout=subprocess.PIPE, text=True) for r in range(n)]
        outs = [p.communicate()[0].strip() for p in ps]
        rates = [float(o.split("once")[1].split("+")[0]) for o in outs if "once" in o]
        print("\n".join(outs))
        print("%d GPU(s)%s: aggregate %.1f GB/s per direction, min %.1f max %.1f" % (n, " bound" if bind else "", sum(rates), min(rates), max(rates)))
