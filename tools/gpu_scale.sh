#!/bin/bash
# usage: gpu_scale.sh N   -- bench.py (default workload c5, strong scaling) at N ranks, as the driver launches it
N=$1
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
  timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
else
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
fi
echo "rc=$?" >> gpurun_out/scale_n$N.err
grep "^rank\|rc=\|Error\|error" gpurun_out/scale_n$N.err | sort | head -12
python - <<PY
import json
for l in open("gpurun_out/scale_n$N.json"):
    if l.startswith("{"):
        d = json.loads(l)
        print({k: d.get(k) for k in ("value", "n_gpus", "ms_per_step", "scaling")}, "with_gather", d.get("with_gather", {}).get("value"), "e2e", d["e2e"]["value"], "frame", d["e2e"].get("frame", {}).get("value"), "hits", d.get("hits_gathered"))
PY
