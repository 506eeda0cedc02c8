#!/bin/bash
# A/B of kernel variants: gpu_ab.sh "<lone-warp variants>" "<c2 tokens>" "<c5 tokens>" "<c3 tokens>"  (empty string skips a step)
mkdir -p gpurun_out
if [ -n "$1" ]; then timeout 300 python tools/lone_warp.py c2 $1 > gpurun_out/lone_warp_c2.log 2>&1; echo "rc=$?" >> gpurun_out/lone_warp_c2.log; tail -12 gpurun_out/lone_warp_c2.log; fi
if [ -n "$2" ]; then timeout 600 python tools/explore.py c2 $2 > gpurun_out/explore_c2.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c2.log; grep -v "parity" gpurun_out/explore_c2.log | cut -c1-400 | tail -24; fi
if [ -n "$3" ]; then timeout 900 python tools/explore.py c5 $3 > gpurun_out/explore_c5.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c5.log; grep -v "^    stats\|parity" gpurun_out/explore_c5.log | tail -8; fi
if [ -n "$4" ]; then timeout 900 python tools/explore.py c3 $4 > gpurun_out/explore_c3.log 2>&1; echo "rc=$?" >> gpurun_out/explore_c3.log; grep -v "^    stats\|parity" gpurun_out/explore_c3.log | tail -8; fi
