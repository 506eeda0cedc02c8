#!/bin/bash
# counted-leaf A/B: GPU tests, then variants with (v6 v1 v7) and without (v8 v9 v10) counted leaves, single-set counted (v12 v11)
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/explore.py c2 v6 v8 v12 v1 v9 > gpurun_out/explore_c2.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c2.log
grep -v "^    stats" gpurun_out/explore_c2.log | tail -12
timeout 300 python tools/lone_warp.py c2 6 8 12 > gpurun_out/lone_warp_c2.log 2>&1; echo "rc=$?" >> gpurun_out/lone_warp_c2.log
tail -8 gpurun_out/lone_warp_c2.log
timeout 900 python tools/explore.py c5 v1 v9 v6 > gpurun_out/explore_c5.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c5.log
grep -v "^    stats" gpurun_out/explore_c5.log | tail -8
timeout 900 python tools/explore.py c3 v7 v10 v11 > gpurun_out/explore_c3.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c3.log
grep -v "^    stats" gpurun_out/explore_c3.log | tail -8
