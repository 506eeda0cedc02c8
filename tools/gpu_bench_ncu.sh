#!/bin/bash
# bench + ncu launch list + one full capture of the traversal kernel (run under gpurun)
mkdir -p gpurun_out
[ -n "$SKIP_BENCH" ] || timeout 600 python bench.py "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline $@"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:trace(Wide)?Kernel" -s 8 -c 2 -o gpurun_out/prof_trace $CMD > gpurun_out/ncu_full.log 2>&1
tail -n 3 gpurun_out/bench.err
cat gpurun_out/bench.json
