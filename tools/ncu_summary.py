"""Turn an .ncu-rep (ncu --set full) into the markdown summary kept under profiles/.
usage: python tools/ncu_summary.py gpurun_out/prof_trace.ncu-rep profiles/rNN_name.md "title" [--traffic profiles/traffic.json key [rays per captured launch]]"""
import csv, io, json, subprocess, sys

METRICS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed.avg.per_cycle_elapsed",
    "smsp__issue_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
    "sm__cycles_active.min", "sm__cycles_active.avg", "sm__cycles_active.max", "sm__cycles_elapsed.max",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts.sum",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "derived__memory_l1_wavefronts_shared_excessive",
]


def main():
    rep, out, title = sys.argv[1], sys.argv[2], sys.argv[3]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    head, units, data = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(head)}
    stall = sorted(n for n in head if n.startswith("smsp__average_warps_issue_stalled") and n.endswith("_per_issue_active.ratio"))
    names = [m for m in METRICS if m in col] + stall
    with open(out, "w") as f:
        f.write("# %s\n\n" % title)
        f.write("| metric | unit | " + " | ".join("launch %d" % i for i in range(len(data))) + " |\n")
        f.write("|---|---|" + "---|" * len(data) + "\n")
        f.write("| kernel | | " + " | ".join(r[col["Kernel Name"]] for r in data) + " |\n")
        for n in names:
            f.write("| %s | %s | %s |\n" % (n, units[col[n]], " | ".join(r[col[n]] for r in data)))
    if "--traffic" in sys.argv:
        k = sys.argv.index("--traffic")
        path, key = sys.argv[k + 1], sys.argv[k + 2]
        try:
            tj = json.load(open(path))
        except Exception:
            tj = {}
        def val(r, n):
            v = float(r[col[n]].replace(",", ""))
            u = units[col[n]].lower()
            return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        per = [val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum") for r in data]
        tj[key] = {"dram_bytes_per_launch": sum(per) / len(per), "launches": len(per), "source": out}
        # executed warp-instructions per launched ray (bench.py's issue roofline): rays of the captured launch = 4th argument
        if len(sys.argv) > k + 3:
            rays = float(sys.argv[k + 3])
            inst = [float(r[col["smsp__inst_executed.sum"]].replace(",", "")) for r in data]
            tj[key]["warp_inst_per_ray"] = sum(inst) / len(inst) / rays
            tj[key]["rays_per_captured_launch"] = rays
            tj[key]["issue_active_per_cycle"] = float(data[0][col["smsp__issue_active.avg.per_cycle_active"]])
            tj[key]["threads_per_inst"] = float(data[0][col["smsp__thread_inst_executed_per_inst_executed.ratio"]])
        json.dump(tj, open(path, "w"), indent=1)
    print(open(out).read())


if __name__ == "__main__":
    main()
