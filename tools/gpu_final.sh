#!/bin/bash
# end-of-round numbers with the library defaults: all GPU tests, smoke, bench, C2-C5 against the reference CUDA tracer
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
for c in c2 c3 c4 c5; do timeout 900 python tools/explore.py $c default > gpurun_out/explore_${c}_default.log 2>&1; echo "rc=$?" >> gpurun_out/explore_${c}_default.log; done
timeout 600 python tools/pipeline_c3.py > gpurun_out/pipeline_c3.log 2>&1; echo "rc=$?" >> gpurun_out/pipeline_c3.log
timeout 900 python bench.py --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "bench c5 rc=$?" >> gpurun_out/bench_c5.err
tail -n 3 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench.err gpurun_out/pipeline_c3.log gpurun_out/bench_c5.err
grep -h "^default\|^ref\|parity" gpurun_out/explore_c?_default.log
cut -c1-400 gpurun_out/bench.json; cut -c1-300 gpurun_out/bench_c5.json
