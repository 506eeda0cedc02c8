#!/bin/bash
# usage: gpu_explore.sh <config> <variant tokens...>   -> gpurun_out/explore_<config>.log
mkdir -p gpurun_out
cfg=$1; shift
timeout 1200 python tools/explore.py $cfg "$@" > gpurun_out/explore_$cfg.log 2>&1; echo "rc=$?" >> gpurun_out/explore_$cfg.log
cat gpurun_out/explore_$cfg.log
