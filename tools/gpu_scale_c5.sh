#!/bin/bash
# usage: gpu_scale_c5.sh N   -- bench.py --workload c5 at N ranks (strong scaling of the 64 Mi-ray diffuse batch)
N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --workload c5 > gpurun_out/scale_c5_n$N.json 2> gpurun_out/scale_c5_n$N.err; echo "rc=$?" >> gpurun_out/scale_c5_n$N.err
grep "^rank" gpurun_out/scale_c5_n$N.err | sort | head -8
cut -c1-200 gpurun_out/scale_c5_n$N.json
