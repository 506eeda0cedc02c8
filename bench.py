#!/usr/bin/env python
"""bench.py -- Mrays/s of the batched BVH trace on BASELINE.json's configs (one JSON line on stdout).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5] [--impl b2rt|reference]

Default workload = c5, the configuration the north_star's target is quoted on (BASELINE.json configs[4]): ONE
64 Mi-ray incoherent diffuse batch on the 910 348-triangle knot (SBVH).  A STEP is one closest-hit trace of that batch
with the rays resident in HBM.  `value` = counted rays of the step / CUDA-event time of the step's trace launches (the
reference's rule, App.cc:194-204: only trace launches are timed; degenerate secondary rays are launched but not
counted, Renderer.cc:221-238).  The batch (2 GiB of rays + 1 GiB of records) is 24 x the 126 MB L2, so nothing
survives from one step to the next and no flush is needed; frame-sized workloads (c1-c4) flush L2 between steps.

N > 1 (torchrun, one process per GPU): STRONG scaling -- rank 0 builds the BVH and broadcasts it (NCCL), the 64 Mi rays
are dealt out in chunks of one 8x8 pixel block (64 primaries x 86 samples), every rank traces its share, time = max
over ranks.  `value` is the trace alone; `with_gather` adds the all-gather of the 16-byte hit records to every step.

Beside it, in the same line:
  e2e            the same batch through b2rt_trace_host_begin/_end with pinned HOST rays in and HOST records out
  e2e.frame      the reference's real frame (Renderer.cc:112-152,242-300): camera -> primary rays -> trace -> count ->
                 diffuse batches generated on the device -> trace -> reconstruct -> pixels to the host
  gpu_reference  the reference's own CUDA kernel (oracle/_ref/libref_gpu.so = kepler_dynamic_fetch.cu built for sm_100a)
                 on the same device arrays and rays: literal 720-warp launch and warps scaled to the SM count
  simd_eff       TRV / IST SIMD efficiency from one untimed instrumented pass (README.md:46-81 columns)
  others         the one-GPU values of c2, c3, c4 with the same reference-kernel ratio (N = 1 only)
  cpu_baseline   the reference's BVH::trace on the host cores (oracle/_ref/ref_cpu), bounded sample (N = 1 only)

--impl reference times the reference's own CPU tracer on the same workload with every host core.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "gpu-ray-tracing_b200"
T_START = time.time()


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# stdout carries exactly ONE line, the JSON result: everything else a library prints there (NCCL's version banner, the reference
# launcher's grid print) goes to stderr.  File descriptor 1 is pointed at stderr for the run; emit() writes to the real stdout.
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def _nbytes(a):
    """bytes of a numpy array or a torch tensor (the broadcast BVH arrays stay on the device under nccl)"""
    return int(a.numel() * a.element_size()) if hasattr(a, "element_size") else int(a.nbytes)


def emit(line):
    _REAL_STDOUT.write(json.dumps(line) + "\n")
    _REAL_STDOUT.flush()


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ---------------------------------------------------------------------------------------------
# clocks during the timed region (NVML)
# ---------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception as e:          # pragma: no cover
            log("clock sampling: NVML unavailable (%s)" % e)
            self._h = None

    def sample(self):
        if self._h is None:
            return
        try:
            nv = self._nv
            self.samples.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
            r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
            for bit, name in self.REASONS.items():
                if (r & bit) and name != "gpu_idle":
                    self.reasons.add(name)
        except Exception:
            pass

    def run(self):
        while not self._stop_evt.is_set():
            self.sample()
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2.0)
        self.sample()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


def visible_to_physical(index):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[index])
        except Exception:
            return index
    return index


# ---------------------------------------------------------------------------------------------
# what both arms say about the workload: identical keys AND values, so the driver's same_config holds
# ---------------------------------------------------------------------------------------------
def bench_config(name, n_gpus):
    wl = importlib.import_module(PKG + ".workloads")
    c = wl.CONFIGS[name]
    strong = bool(c.get("total_rays"))
    return {
        "workload": wl.describe(name),
        "step": wl.describe_step(name),
        "bvh": "SBVH (split alpha 1e-5, leaves of 1..8 triangles), BVH replicated on every GPU",
        "parallelism": ("rays sharded over %d GPU(s) in 8x8-pixel-block chunks" % n_gpus) if strong
                       else ("one frame per GPU x %d" % n_gpus),
        "l2": "batch is 24x the L2: nothing survives between steps" if strong else "flushed between timed steps (256 MiB write)",
    }


# ---------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's own BVH::trace (oracle/_ref/ref_cpu)
# ---------------------------------------------------------------------------------------------
def _workload_obj(name, tmpdir):
    """OBJ + camera signature of a workload.  The geometry comes from libb2rt_scenes.so (the seeded scene generators, a
    host-only library of their own): libb2rt.so is NOT loaded on this path."""
    pkg = importlib.import_module(PKG)
    wl = importlib.import_module(PKG + ".workloads")
    c = wl.CONFIGS[name]
    sl = pkg.scenes_lib()
    h = C.c_void_p()
    if sl.b2rt_scene_generate(pkg.host.SCENE_KINDS[c["kind"]], c["tris"], C.c_uint64(c["seed"]), C.byref(h)) != 0:
        raise RuntimeError("scene generation failed")
    os.makedirs(tmpdir, exist_ok=True)
    path = os.path.join(tmpdir, "%s-%d-%d.obj" % (c["kind"], c["tris"], c["seed"]))
    if not os.path.exists(path):
        tmp = path + ".%d.tmp" % os.getpid()
        if sl.b2rt_scene_write_obj(h, tmp.encode()) != 0:
            raise RuntimeError("cannot write " + tmp)
        os.replace(tmp, path)
    cam = (C.c_float * 12)()
    sl.b2rt_scene_camera(h, cam)
    sl.b2rt_scene_free(h)
    camera = {"pos": list(cam[0:3]), "forward": list(cam[3:6]), "up": list(cam[6:9]), "fov": cam[9], "near": cam[10], "far": cam[11]}
    sig = pkg.encode_camera_signature(camera).strip('",')
    return path, sig, camera, c


class ReferenceCpu:
    """One run of oracle/_ref/ref_cpu: imports the OBJ, builds the SBVH with the reference's SplitBVHBuilder (single
    threaded; 73 s for the 910 K-triangle knot), then times BVH::trace (BVH.cc:67-163) `repeats` times on `procs` forked
    workers.  start() returns at once; with hold=True the trace waits for release(), so a caller can overlap the build
    with its own (GPU) work and still time the trace on idle cores."""

    def __init__(self, name, repeats, procs, budget_s=0.0, hold=False, tmpdir="/tmp/b2rt_bench"):
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as O
        self.O, self.name, self.repeats, self.procs = O, name, repeats, procs
        self.proc = None
        self.sample = ""
        if not O.have_ref_cpu():
            return
        os.makedirs(tmpdir, exist_ok=True)
        obj, sig, camera, c = _workload_obj(name, tmpdir)
        w, h = c["size"]
        args = [O.REF_CPU, "--obj", obj, "--camera-sig", sig, "--size", w, h, "--procs", procs, "--repeat", repeats]
        self.secondary = c["ray_type"] != "primary"
        self.with_primary = bool(c.get("with_primary")) or not self.secondary
        if self.secondary:
            if c["ray_type"] == "ao":
                raise NotImplementedError("reference arm covers primary / diffuse workloads")
            args += ["--trace-secondary", "--secondary-closest", "--samples", c["samples"], "--max-dist", camera["far"]]
            if c.get("total_rays"):
                # the measured prefix of the big batch: the first 2 Mi-ray secondary batch (RayGen's batch size,
                # Renderer.cc:46) generated from every k-th 8x8 pixel block of the frame, so that it covers the whole
                # frame -- 23 872 primaries x 86 samples = 2 052 992 rays, no extrapolation
                blocks = (w * h) // 64
                want_blocks = max(1, (1 << 21) // c["samples"] // 64)
                stride = -(-blocks // want_blocks)
                args += ["--prim-stride", stride]
                self.sample = "first 2 Mi-ray diffuse batch from every %d-th 8x8 pixel block of the frame" % stride
        else:
            args += ["--results-out", os.path.join(tmpdir, "ref_primary.%d.results" % os.getpid())]
        if budget_s and not c.get("total_rays"):
            args += ["--budget-s", budget_s]
            self.sample = "every k-th 8x8 pixel block of the frame, sized for a %.0f s budget" % budget_s
        self.go = None
        if hold:
            self.go = os.path.join(tmpdir, "go.%d" % os.getpid())
            if os.path.exists(self.go):
                os.remove(self.go)
            args += ["--go-file", self.go]
        self.t0 = time.time()
        self.proc = subprocess.Popen([str(a) for a in args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)

    def release(self):
        if self.go:
            open(self.go, "w").close()

    def result(self):
        if self.proc is None:
            return None
        self.release()
        out, err = self.proc.communicate()
        if self.go and os.path.exists(self.go):
            os.remove(self.go)
        if self.proc.returncode != 0:
            raise RuntimeError("ref_cpu failed: " + err[-400:])
        prim, sec = [], []
        for line in out.splitlines():
            line = line.strip()
            if line.startswith("{"):
                d = json.loads(line)
                (sec if d.get("stage") == "secondary" else prim).append(d)
        steps = []
        for i in range(self.repeats):
            rays, secs = 0, 0.0
            if self.with_primary:
                rays, secs = prim[i]["rays"], prim[i]["trace_s"]
            if sec:
                rays += sec[i]["counted"]
                secs += sec[i]["trace_s"]
            steps.append({"rays": rays, "seconds": secs})
        build = [l for l in err.splitlines() if "build" in l]
        return {"steps": steps, "wall": time.time() - self.t0, "kind": "reference", "sample": self.sample,
                "build": build[0].strip() if build else ""}


def run_port_cpu(name, repeats, procs):
    """Fallback when oracle/_ref/ref_cpu is absent: the oracle's restatement of BVH::trace
    (b2o_trace_mt) over forked workers, on the product-built BVH (primary rays only)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_lib as O
    from concurrent.futures import ProcessPoolExecutor
    b2rt = importlib.import_module(PKG)
    wl = importlib.import_module(PKG + ".workloads")
    c = wl.CONFIGS[name]
    scene, bvh = wl.build_scene_and_bvh(name, log=log)
    w, h = c["size"]
    origin, m = b2rt.camera_matrices(scene.camera, w, h)
    i2p, _ = O.pixel_table(w, h)
    rays, _, _ = O.raygen_primary(origin, m, w, h, scene.camera["far"], i2p)
    arrays = {k: np.array(v) for k, v in bvh.arrays().items() if k != "layout"}
    sc = {"verts": scene.verts, "tris": scene.tris}
    global _PORT_STATE
    _PORT_STATE = (arrays, sc, rays)
    steps = []
    n = len(rays)
    bounds = [(n * p // procs, n * (p + 1) // procs) for p in range(procs)]
    import multiprocessing as mp
    with ProcessPoolExecutor(procs, mp_context=mp.get_context("fork")) as ex:
        for _ in range(repeats):
            t0 = time.time()
            list(ex.map(_port_worker, bounds))
            steps.append({"rays": n, "seconds": time.time() - t0})
    return {"steps": steps, "wall": sum(s["seconds"] for s in steps), "kind": "port", "sample": "primary rays of the frame", "build": ""}


_PORT_STATE = None


def _port_worker(b):
    import oracle_lib as O
    arrays, sc, rays = _PORT_STATE
    O.trace_mt(arrays, sc, rays[b[0]:b[1]])
    return b[1] - b[0]


def reference_arm(args):
    """bench.py --impl reference: rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    procs = args.cores if args.cores > 0 else host_cores()
    reps = args.warmup + args.steps
    ref = ReferenceCpu(args.workload, reps, procs, budget_s=120.0)
    r = ref.result()
    if r is None:
        r = run_port_cpu(args.workload, min(reps, 8), procs)
    steps = r["steps"][min(args.warmup, len(r["steps"]) - 1):]
    rays = sum(s["rays"] for s in steps)
    secs = sum(s["seconds"] for s in steps)
    value = rays / secs * 1e-6
    wl = importlib.import_module(PKG + ".workloads")
    strong = bool(wl.CONFIGS[args.workload].get("total_rays"))
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / len(steps) * 1e3,
        "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(args.workload, args.gpus),
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": procs, "kind": r["kind"],
                         "sample": "%d counted rays per step (%s), %d timed steps; reference SplitBVHBuilder tree (%s)"
                                   % (steps[0]["rays"], r["sample"] or "the whole frame", len(steps), r["build"])},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------------------------
# the product arm
# ---------------------------------------------------------------------------------------------
def algorithmic_bytes(work):
    """B = 32 (ray) + 16 (result) + 64 N_n + 48 N_t + 16 N_l + 4 h, summed over a batch."""
    return 48 * work["rays"] + 64 * work["nodes"] + 48 * work["tris"] + 16 * work["leaves"] + 4 * work["hits"]


def algorithmic_fp32(work):
    """F = 36 N_n + 8 N_t + 7 N_u + 8 N_v + 9 per ray (FMA = 1)."""
    return 36 * work["nodes"] + 8 * work["tris"] + 7 * work["ustage"] + 8 * work["vstage"] + 9 * work["rays"]


def time_reference_kernel(frame, sm_count, oracle, reps=5):
    """The reference's CUDA tracer on the frame's batches (same device arrays, same rays).  Returns
    {label: ms per step} for the literal 720-warp launch (CudaTracer.cc:156) and for warps scaled to the SM count, and the
    records of the last scaled run (for the agreement figure)."""
    import numpy as np
    import torch
    ref = oracle.RefGpuTracer()
    ref.set_bvh(frame.cuda_bvh.m_nodes, frame.cuda_bvh.m_triWoop, frame.cuda_bvh.m_triIndex)
    out = {}
    res = [torch.zeros_like(b.m_results) for b, _ in frame.batches]
    for label, warps in (("literal_720_warps", 720), ("scaled_32_per_sm", sm_count * 32), ("scaled_40_per_sm", sm_count * 40),
                         ("scaled_48_per_sm", sm_count * 48)):
        def step():
            return sum(ref.trace(b.getRayBuffer_dev(), r, not b.getNeedClosestHit(), warps) for (b, _), r in zip(frame.batches, res))
        step()
        step()
        out[label] = float(np.median([step() for _ in range(reps)]))
    return out, res


def measure_frame(frame, flush, steps, warmup):
    """ms per step (summed CUDA-event time of the frame's trace launches), L2 flushed before every step when asked."""
    for _ in range(warmup):
        if flush is not None:
            flush.add_(1)
        frame.trace_all()
    ms = []
    for _ in range(steps):
        if flush is not None:
            flush.add_(1)
        ms.append(frame.trace_all())
    return ms


def simd_efficiency(frame):
    """One untimed pass of the instrumented kernel over every batch (B2RT_TRACE_STATS)."""
    tot = {"node_iters": 0, "node_lanes": 0, "tri_iters": 0, "tri_lanes": 0, "fetches": 0}
    frame.tracer.collectStats = True
    try:
        for rb, _ in frame.batches:
            frame.tracer.traceBatch(rb)
            st = frame.tracer.getStats()
            for k in tot:
                tot[k] += st[k]
    finally:
        frame.tracer.collectStats = False
    return {"trv_pct": 100.0 * tot["node_lanes"] / max(1, 32 * tot["node_iters"]),
            "ist_pct": 100.0 * tot["tri_lanes"] / max(1, 32 * tot["tri_iters"]),
            "warp_node_steps": tot["node_iters"], "warp_triangle_steps": tot["tri_iters"], "warp_ray_fetches": tot["fetches"],
            "how": "mean active lanes / 32 over node-loop and triangle-loop iterations of the instrumented kernel (README.md:46-81)"}


def id_agreement(frame, ref_res):
    import numpy as np
    same = total = 0
    worst = 0.0
    for (b, _), r in zip(frame.batches, ref_res):
        n = b.getSize()
        mine = b.getResultBuffer_dev()
        if not b.getNeedClosestHit():
            same += int(((mine[:, 0] >= 0) == (r[:n, 0] >= 0)).sum().item())
            total += n
            continue
        eq = mine[:, 0] == r[:n, 0]
        same += int(eq.sum().item())
        total += n
        hit = eq & (mine[:, 0] >= 0)
        if bool(hit.any()):
            tm = mine[:, 1][hit].view(__import__("torch").float32)
            tr = r[:n, 1][hit].view(__import__("torch").float32)
            worst = max(worst, float(((tm - tr).abs() / tr.abs()).max().item()))
    return same / max(1, total), worst


def product_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    b2rt = importlib.import_module(PKG)
    wl = importlib.import_module(PKG + ".workloads")
    mg = importlib.import_module(PKG + ".multigpu")
    binding = mg.bind_to_local_cores(local_rank, ws)          # before any pinned allocation: first-touch decides the NUMA node
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if ws > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    name = args.workload
    cfg = wl.CONFIGS[name]
    strong = bool(cfg.get("total_rays"))

    # ---- CPU baseline: its single-threaded BVH build runs in the background while the GPU works; its timed trace is
    #      released only after every GPU measurement is done (N = 1, rank 0 only)
    cpu_ref = None
    if ws == 1 and not args.no_cpu_baseline:
        try:
            cpu_ref = ReferenceCpu(name, 2, host_cores(), budget_s=20.0, hold=True)
        except Exception as e:
            log("cpu baseline not started: %s" % e)

    ctx = b2rt.Context.get(local_rank)
    if args.variant >= 0 or args.top_nodes >= 0 or args.tune:
        ctx.set_trace_config(variant=args.variant, top_nodes=args.top_nodes, tune=args.tune)

    # ---- scene + BVH: rank 0 builds, everybody gets the arrays (NCCL broadcast) ----
    scene = b2rt.Scene.generate(cfg["kind"], cfg["tris"], cfg["seed"])          # cheap and deterministic on every rank
    arrays = None
    if rank == 0:
        _, bvh = wl.build_scene_and_bvh(name, log=log)
        arrays = {k: (np.array(v) if k != "layout" else v) for k, v in bvh.arrays().items()}
    arrays = mg.broadcast_bvh(arrays, device)
    shard = (rank, ws) if strong else (0, 1)
    frame = wl.Frame(name, scene, arrays, device, ctx, shard=shard)
    counted = frame.counted_rays()
    launched = frame.launched_rays()
    log("rank %d: %d batches, %d rays launched, %d counted" % (rank, len(frame.batches), launched, counted))

    # ---- algorithmic work (canonical counters), once, untimed ----
    work = {}
    for rb, _ in frame.batches:
        wk = frame.tracer.countWork(rb)
        for k, v in wk.items():
            work[k] = max(work.get(k, 0), v) if k in ("max_stack", "max_chain") else work.get(k, 0) + v
    bytes_step = algorithmic_bytes(work)
    fp32_step = algorithmic_fp32(work)

    flush = None if strong else torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=device)     # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if ws > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident throughput -----------------------------------------------------------
    measure_frame(frame, flush, 0, args.warmup)
    sampler = ClockSampler(visible_to_physical(local_rank))
    launches0 = ctx.launch_count()
    barrier()
    sampler.start()
    t_wall0 = time.time()
    ms_steps = measure_frame(frame, flush, args.steps, 0)          # CUDA events on the launching stream, per trace launch
    barrier()
    t_wall = time.time() - t_wall0
    gpu_launches = ctx.launch_count() - launches0
    ms_total = float(sum(ms_steps))

    # ---- the same step with the hit records all-gathered (N > 1, sharded workloads) ----------------
    gather_ms = 0.0
    hits_gathered = None
    last_rb = frame.batches[-1][0]
    if ws > 1 and strong:
        g_steps = max(3, min(args.steps, 10))
        per = mg.max_over_ranks(last_rb.getSize(), device)
        pad = torch.zeros((per, 4), dtype=torch.int32, device=device)
        gathered = torch.empty((ws * per, 4), dtype=torch.int32, device=device)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(2 + g_steps):
            if it == 2:
                barrier()
                ev0.record()
            frame.tracer.traceBatchAsync(last_rb)
            pad[: last_rb.getSize()].copy_(last_rb.getResultBuffer_dev(), non_blocking=True)
            dist.all_gather_into_tensor(gathered, pad)
        ev1.record()
        barrier()
        gather_ms = ev0.elapsed_time(ev1) / g_steps
        hits_gathered = int((gathered[:, 0] >= 0).sum().item())       # pad rows are id 0 ... count only real rows below
        # pad rows carry id 0 (zeros): subtract them -- every rank knows the slice sizes
        sizes = mg.all_sizes(last_rb.getSize(), device)
        hits_gathered -= sum(per - s for s in sizes)
        del gathered, pad
    elif not strong and ws > 1:
        gathered = mg.gather_results(last_rb.getResultBuffer_dev(), last_rb.getSize() * ws, device)
        hits_gathered = int((gathered[:, 0] >= 0).sum().item())
        del gathered
    else:
        hits_gathered = int((last_rb.getResultBuffer_dev()[:, 0] >= 0).sum().item())

    # ---- end to end through the C ABI with HOST buffers ------------------------------------------
    host_rays = [torch.empty((rb.getSize(), 8), dtype=torch.float32).pin_memory() for rb, _ in frame.batches]
    host_res = [torch.empty((rb.getSize(), 4), dtype=torch.int32).pin_memory() for rb, _ in frame.batches]
    for hr, (rb, _) in zip(host_rays, frame.batches):
        hr.copy_(rb.getRayBuffer_dev())
    h2d = sum(hr.numel() * 4 for hr in host_rays)
    d2h = sum(hr.numel() * 4 for hr in host_res)

    def e2e_step():
        # the frame's batches are independent host buffers: begin them all, wait once (b2rt_trace_host_begin / _end)
        for hr, hs, (rb, _) in zip(host_rays, host_res, frame.batches):
            frame.tracer.traceHostBegin(hr, hs, anyHit=not rb.getNeedClosestHit())
        frame.tracer.traceHostEnd()

    for _ in range(2):
        e2e_step()
    e2e_steps = max(3, min(args.steps, 10 if strong else 50))
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    ev0.record()
    for _ in range(e2e_steps):
        e2e_step()
    ev1.record()
    torch.cuda.synchronize()
    e2e_wall = time.time() - t0
    e2e_ms = max(ev0.elapsed_time(ev1), e2e_wall * 1e3)         # host-synchronous calls: wall clock is the honest one
    same = float((host_res[-1][:, 0] == last_rb.getResultBuffer_dev()[:, 0].cpu()).float().mean())
    assert same >= 0.9999, "b2rt_trace_host and b2rt_trace disagree (id agreement %.6f)" % same
    del host_rays, host_res

    # ---- the reference's real frame: everything born on the device, pixels back (Renderer.cc:112-152,242-300) ----
    frame_e2e = None
    if cfg["ray_type"] == "diffuse":
        frame_one = wl.render_frame_e2e(name, scene, frame.cuda_bvh, ctx, device, steps=max(2, min(args.steps, 5)), shard=shard, n_streams=1)
        frame_e2e = wl.render_frame_e2e(name, scene, frame.cuda_bvh, ctx, device, steps=max(2, min(args.steps, 5)), shard=shard, n_streams=2)
        # same image either way (pixels are written by disjoint reconstruct launches)
        assert bool((frame_one.pop("pixels") == frame_e2e.pop("pixels")).all()), "frame: two-stream order changed the image"
        frame_e2e["ms_per_step_one_stream"] = frame_one["ms_per_step"]
    sampler.stop()

    # ---- reduce over ranks ----------------------------------------------------------------------
    red = mg.reduce_stats([ms_total, e2e_ms, gather_ms, frame_e2e["ms_per_step"] if frame_e2e else 0.0], device, op="max")
    ms_total_max, e2e_ms_max, gather_ms_max, frame_ms_max = red
    tot = mg.reduce_stats([counted, launched, gpu_launches, bytes_step, fp32_step, h2d, d2h,
                           frame_e2e["counted"] if frame_e2e else 0], device, op="sum")
    counted_all, launched_all, launches_all, bytes_all, fp32_all, h2d_all, d2h_all, frame_counted_all = tot

    if rank != 0:
        if ws > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    value = counted_all * args.steps / (ms_total_max * 1e-3) * 1e-6
    e2e_value = counted_all * e2e_steps / (e2e_ms_max * 1e-3) * 1e-6
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except Exception:
        pass
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    n_launch_step = len(frame.batches)
    avg_launch_ms = ms_total_max / (args.steps * n_launch_step)
    achieved = (bytes_all / ws / n_launch_step) / (avg_launch_ms * 1e-3) * 1e-9        # per GPU, per launch
    prof = {}
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            prof = json.load(f).get(name, {})
    except Exception:
        pass
    traffic = prof.get("dram_bytes_per_launch")
    # which kernel the library's default policy launches for these batches (trace.cu:launchTrace)
    wide = (args.variant >= 100) or (args.variant < 0 and float(_nbytes(arrays["nodes"]) + _nbytes(arrays["woop"])) > float(48 << 20))
    clocks = sampler.summary()
    sm_hz = (clocks["sm_mhz"] or peaks.get("sm_max_mhz", 1965.0)) * 1e6
    sm_count = ctx.sm_count()
    issue_peak = sm_count * 4 * sm_hz                      # warp-instructions per second: 4 schedulers per SM, one issue per clock
    # executed warp-instructions per launched ray from the committed ncu capture of this kernel on this workload
    # (profiles/traffic.json: smsp__inst_executed.sum / rays of the captured launch), times the rays of one launch here
    inst_per_ray = prof.get("warp_inst_per_ray")
    rays_per_launch = launched / n_launch_step
    issue = None
    if inst_per_ray:
        inst_rate = inst_per_ray * rays_per_launch / (avg_launch_ms * 1e-3)
        issue = {"achieved": inst_rate * 1e-9, "peak": issue_peak * 1e-9, "frac": inst_rate / issue_peak}
    fp32_rate = fp32_all / ws / (ms_total_max / args.steps * 1e-3)
    hbm = {"bound": "hbm", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs, "peak_source": peak_src,
           "algorithmic_bytes_per_launch": bytes_all / ws / n_launch_step,
           "note": "SURVEY 8(d) algorithmic bytes count every node / triangle fetch; the BVH is L1/L2 resident, so this exceeds "
                   "the DRAM ceiling by construction and does not bound the kernel -- `traffic` is what really crosses the HBM interface"}
    roofline = {
        "bound": "issue", "unit": "G warp-inst/s",
        "achieved": issue["achieved"] if issue else None, "peak": issue_peak * 1e-9, "frac": issue["frac"] if issue else None,
        "traffic": traffic, "kernel": "traceWideKernel" if wide else "traceKernel", "avg_launch_ms": avg_launch_ms,
        "how": "executed warp-instructions per ray (ncu smsp__inst_executed.sum of this kernel on this workload, profiles/traffic.json) x rays per "
               "launch / live CUDA-event launch time, against 4 schedulers x %d SMs x the SM clock sampled during the timed region" % sm_count
               + ("" if ws == 1 else "; the per-ray instruction count is the one-GPU capture's (whole batch, block-claim kernel): a shard below 12 Mi rays runs "
                                     "the kernel without block claims, which executes a few per cent more per ray"),
        "fp32": {"inst_per_s": fp32_rate, "peak_inst_per_s": sm_count * 128 * sm_hz, "frac": fp32_rate / (sm_count * 128 * sm_hz),
                 "how": "SURVEY 8(d) algorithmic FP32 instructions (canonical traversal) against 128 lanes per SM"},
        "hbm": hbm,
    }
    line = {
        "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": ws, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total_max / args.steps, "higher_is_better": True,
        "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(name, ws),
        "detail": {"rays_per_step_per_gpu": launched, "counted_rays_per_step": int(counted_all), "launched_rays_per_step": int(launched_all),
                   "batches_per_step": n_launch_step,
                   "bvh": "%d nodes, %d Woop rows" % (_nbytes(arrays["nodes"]) // 64, _nbytes(arrays["woop"]) // 16)},
        "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all),
                "steps": e2e_steps, "ms_per_step": e2e_ms_max / e2e_steps,
                "api": "b2rt_trace_host_begin per batch + b2rt_trace_host_end per step (pinned host rays in, host hit records out)",
                "host_binding": binding},
        "gpu_launches": int(launches_all),
        "wall_s_timed_region": t_wall,
        "clocks": clocks,
        "roofline": roofline,
        "work_per_ray": {k: work[k] / max(1, work["rays"]) for k in ("nodes", "tris", "leaves", "ustage", "vstage", "hits")},
        "hits_gathered": hits_gathered, "host_vs_device_id_agreement": same,
    }
    if frame_e2e:
        line["e2e"]["frame"] = {
            "value": frame_counted_all / (frame_ms_max * 1e-3) * 1e-6, "unit": "Mrays/s", "ms_per_step": frame_ms_max,
            "h2d_bytes_per_step": frame_e2e["h2d"] * ws, "d2h_bytes_per_step": frame_e2e["d2h"] * ws, "launches_per_step": frame_e2e["launches"],
            "api": frame_e2e["api"],
            "ms_per_step_one_stream_rank0": frame_e2e["ms_per_step_one_stream"]}
    if ws > 1 and strong:
        t_g = gather_ms_max
        line["with_gather"] = {"value": counted_all / (t_g * 1e-3) * 1e-6, "unit": "Mrays/s", "ms_per_step": t_g,
                               "bytes_gathered_per_step": int(launched_all) * 16,
                               "what": "trace + ncclAllGather of every rank's 16-byte hit records, CUDA events around both, max over ranks"}

    # ---- SIMD efficiency (README.md:46-81 TRV / IST), one untimed instrumented pass ----
    try:
        line["simd_eff"] = simd_efficiency(frame)
    except Exception as e:
        line["simd_eff"] = {"error": str(e)}

    # ---- the reference's CUDA kernel on the same arrays and rays (N = 1) ----
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    oracle = None
    if ws == 1 and not args.no_gpu_reference:
        try:
            import oracle_lib as oracle
            if not oracle.have_ref_gpu():
                oracle = None
        except Exception:
            oracle = None
    if oracle is not None:
        try:
            frame.trace_all()
            ref_ms, ref_res = time_reference_kernel(frame, sm_count, oracle)
            agree, worst = id_agreement(frame, ref_res)
            best = min(v for k, v in ref_ms.items() if k.startswith("scaled"))
            mine_ms = ms_total_max / args.steps
            line["gpu_reference"] = {
                "kernel": "kepler_dynamic_fetch.cu (the reference's tracer, built for sm_100a: oracle/_ref/libref_gpu.so), same device arrays and rays, same process",
                "mrays_per_s": {k: counted / v * 1e-3 for k, v in ref_ms.items()},
                "ratio_vs_scaled": best / mine_ms, "ratio_vs_literal": ref_ms["literal_720_warps"] / mine_ms,
                "id_agreement": agree, "max_rel_t_error_on_equal_ids": worst}
            del ref_res
        except Exception as e:
            line["gpu_reference"] = {"error": str(e)}

    # ---- the other one-GPU configurations, with the same ratio (N = 1) ----
    if ws == 1 and args.others:
        del frame
        torch.cuda.empty_cache()
        line["others"] = {}
        for other in args.others.split(","):
            if other == name or other not in wl.CONFIGS:
                continue
            if time.time() - T_START > args.others_deadline:
                line["others"][other] = {"skipped": "time budget of the default run (%d s) used up" % args.others_deadline}
                continue
            try:
                line["others"][other] = measure_other(other, ctx, device, sm_count, oracle, b2rt, wl)
            except Exception as e:
                line["others"][other] = {"error": str(e)[:200]}
            torch.cuda.empty_cache()

    # ---- CPU baseline beside it (rank 0, N = 1 only): released now, so that its workers have the cores to themselves ----
    if cpu_ref is not None:
        procs = host_cores()
        try:
            r = cpu_ref.result()
            if r is None:
                r = run_port_cpu(name, 2, procs)
            s = r["steps"][-1]
            line["cpu_baseline"] = {"value": s["rays"] / s["seconds"] * 1e-6, "unit": "Mrays/s", "cores": procs, "kind": r["kind"],
                                    "sample": "%d counted rays (%s) after one warm-up pass; BVH::trace over %d forked workers; reference SplitBVHBuilder tree (%s)"
                                              % (s["rays"], r["sample"] or "the whole frame", procs, r["build"])}
        except Exception as e:          # the baseline must never take the bench line down with it
            line["cpu_baseline"] = {"value": None, "unit": "Mrays/s", "cores": procs, "kind": "reference", "sample": "failed: %s" % e}
    line["bench_wall_s"] = time.time() - T_START
    emit(line)
    if ws > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def measure_other(name, ctx, device, sm_count, oracle, b2rt, wl, steps=5, warmup=3):
    import torch
    t0 = time.time()
    scene, bvh = wl.build_scene_and_bvh(name, log=log)
    frame = wl.Frame(name, scene, bvh.arrays(), device, ctx)
    counted = frame.counted_rays()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=device)
    ms = measure_frame(frame, flush, steps, warmup)
    per_step = sum(ms) / len(ms)
    out = {"workload": wl.describe(name), "value": counted / per_step * 1e-3, "unit": "Mrays/s", "ms_per_step": per_step,
           "counted_rays_per_step": counted, "batches_per_step": len(frame.batches), "simd_eff": None}
    try:
        se = simd_efficiency(frame)
        out["simd_eff"] = {"trv_pct": se["trv_pct"], "ist_pct": se["ist_pct"]}
    except Exception as e:
        out["simd_eff"] = {"error": str(e)[:100]}
    if oracle is not None:
        frame.trace_all()
        ref_ms, ref_res = time_reference_kernel(frame, sm_count, oracle, reps=3)
        agree, worst = id_agreement(frame, ref_res)
        best = min(v for k, v in ref_ms.items() if k.startswith("scaled"))
        out["gpu_reference"] = {"mrays_per_s_scaled": counted / best * 1e-3, "mrays_per_s_literal": counted / ref_ms["literal_720_warps"] * 1e-3,
                                "ratio_vs_scaled": best / per_step, "ratio_vs_literal": ref_ms["literal_720_warps"] / per_step,
                                "id_agreement": agree, "max_rel_t_error_on_equal_ids": worst}
    out["setup_s"] = time.time() - t0
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="c5")
    ap.add_argument("--impl", default="b2rt", choices=["b2rt", "reference"])
    ap.add_argument("--cores", type=int, default=0, help="reference arm: host cores to use (0 = all; 1 = the CPU-1 row of BASELINE.md)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-reference", action="store_true")
    ap.add_argument("--others", default="c2,c3,c4", help="one-GPU configurations measured beside the headline (N = 1); '' = none")
    ap.add_argument("--others-deadline", type=int, default=170, help="do not start another 'others' configuration after this many seconds")
    ap.add_argument("--variant", type=int, default=-1, help="traversal kernel variant (b2rt_trace_config.variant); -1 = library default")
    ap.add_argument("--top-nodes", type=int, default=-1)
    ap.add_argument("--tune", type=int, default=0, help="b2rt_trace_config.tune bits (A/B switches)")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    args.warmup = max(args.warmup, 3)
    return product_arm(args)


if __name__ == "__main__":
    sys.exit(main())
