#!/bin/bash
# the no-allocation entry point alone in a fresh process, then the whole GPU suite (-x), then the bench line
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q ) > gpurun_out/pytest_no_alloc.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_no_alloc.log
tail -n 8 gpurun_out/pytest_no_alloc.log
( time timeout 2400 python -m pytest tests -m gpu -x -q -s ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -v "^{'label'\|^C5 whole" gpurun_out/pytest_gpu.log | tail -n 30
grep "^{'label'\|^C5 whole" gpurun_out/pytest_gpu.log | cut -c1-330


