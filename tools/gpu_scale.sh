#!/bin/bash
# usage: gpu_scale.sh N   -- bench.py at N ranks (torchrun), C2 weak scaling and C5 strong scaling
N=$1
mkdir -p gpurun_out
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
run --steps 100 --warmup 5 > gpurun_out/scale_c2_n$N.json 2> gpurun_out/scale_c2_n$N.err; echo "rc=$?" >> gpurun_out/scale_c2_n$N.err
run --steps 10 --warmup 3 --workload c5 > gpurun_out/scale_c5_n$N.json 2> gpurun_out/scale_c5_n$N.err; echo "rc=$?" >> gpurun_out/scale_c5_n$N.err
tail -n 3 gpurun_out/scale_c2_n$N.err gpurun_out/scale_c5_n$N.err
cat gpurun_out/scale_c2_n$N.json gpurun_out/scale_c5_n$N.json | cut -c1-400
