#!/bin/bash
# round 2, second GPU call: ring fetch A/B on C5, GPU tests, the new default bench line, one full ncu capture of the C5 kernel
mkdir -p gpurun_out
timeout 900 python tools/explore.py c5 v108 v114 default > gpurun_out/explore_c5_ring.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c5_ring.log
cat gpurun_out/explore_c5_ring.log
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 15 gpurun_out/pytest_gpu.log
( time timeout 900 python bench.py ) > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench rc=$?" >> gpurun_out/bench_c5.err
tail -n 12 gpurun_out/bench_c5.err; cat gpurun_out/bench_c5.json
timeout 300 python tools/one_launch.py c5 > gpurun_out/one_launch_c5.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:traceWideKernel" --launch-skip 1 --launch-count 1 -o gpurun_out/r02_wide_c5 python tools/one_launch.py c5 > gpurun_out/ncu_full_c5.log 2>&1
tail -n 3 gpurun_out/one_launch_c5.log gpurun_out/ncu_full_c5.log
