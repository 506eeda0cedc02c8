#!/bin/bash
# round 2, first GPU call: GPU tests after the collapse / margin / fetch changes, then A/B of the new switches on C5 and C4
mkdir -p gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 5 gpurun_out/pytest_gpu.log
# tune bits: 1 no streaming, 2 no refill, 4 greedy collapse, 8 L2 window
timeout 1200 python tools/explore.py c5 v108 v108,x7 v108,x4 v108,x1 v108,x2 v108,x8 v1 v1,x1 > gpurun_out/explore_c5.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c5.log
cat gpurun_out/explore_c5.log
timeout 1200 python tools/explore.py c5 --sorted v108 v1 > gpurun_out/explore_c5_sorted.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c5_sorted.log
cat gpurun_out/explore_c5_sorted.log
timeout 1500 python tools/explore.py c4 v108 v108,x7 v108,x4 v108,x1 > gpurun_out/explore_c4.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c4.log
cat gpurun_out/explore_c4.log
timeout 900 python tools/explore.py c3 default v108,x7 > gpurun_out/explore_c3.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c3.log
cat gpurun_out/explore_c3.log
timeout 600 python tools/explore.py c2 default v8,x1 > gpurun_out/explore_c2.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c2.log
cat gpurun_out/explore_c2.log
