#!/bin/bash
# round-2 verification pass: whole GPU suite, smoke, both bench arms, ncu launch list of the default bench command
mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
tail -n 2 gpurun_out/bench.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?" >> gpurun_out/bench_ref.err
tail -n 1 gpurun_out/bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c5.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-gpu-reference --others '' > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?" >> gpurun_out/ncu_bench.log
tail -n 2 gpurun_out/ncu_bench.log
wc -c gpurun_out/bench.json gpurun_out/bench_ref.json gpurun_out/launches_c5.csv
