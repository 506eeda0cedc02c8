#!/bin/bash
# one full ncu capture of the default trace kernel on a workload (after the same program ran clean without ncu)
W=${1:-c5}
mkdir -p gpurun_out
timeout 300 python tools/one_launch.py $W > gpurun_out/one_launch_$W.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:trace(Wide)?Kernel" --launch-skip 1 --launch-count 1 -f -o gpurun_out/r02_trace_$W python tools/one_launch.py $W > gpurun_out/ncu_full_$W.log 2>&1
tail -n 2 gpurun_out/one_launch_$W.log gpurun_out/ncu_full_$W.log
