#!/bin/bash
# full GPU pass: all GPU tests, smoke, bench (both arms), sorted-AO pipeline, aux kernels
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 600 python tools/pipeline_c3.py > gpurun_out/pipeline_c3.log 2>&1; echo "rc=$?" >> gpurun_out/pipeline_c3.log
timeout 500 python tools/bench_aux.py > gpurun_out/bench_aux.log 2>&1; echo "rc=$?" >> gpurun_out/bench_aux.log
tail -n 4 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench.err gpurun_out/pipeline_c3.log
grep -i sort gpurun_out/bench_aux.log
cat gpurun_out/bench.json
