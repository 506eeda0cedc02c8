#!/bin/bash
# full ncu capture of every trace launch of ONE step of a frame-sized workload (c1: 1 launch, c2: 2), after a clean run
W=${1:-c2}; N=${2:-2}
mkdir -p gpurun_out
timeout 300 python tools/one_launch.py $W > gpurun_out/one_launch_$W.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:trace(Wide)?Kernel" --launch-skip $N --launch-count $N -f -o gpurun_out/r02_trace_$W python tools/one_launch.py $W > gpurun_out/ncu_full_$W.log 2>&1
tail -n 2 gpurun_out/one_launch_$W.log gpurun_out/ncu_full_$W.log
