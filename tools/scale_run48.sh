#!/bin/bash
# 4- and 8-GPU runs of bench.py on one 8-GPU box (final peer exchange)
mkdir -p gpurun_out
for n in 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale2_$n.json
  python - <<PY
import json
d=json.load(open("gpurun_out/scale2_$n.json"))
print("$n", round(d["value"]/1e9,1), round(d["ms_per_step"],4), {k:round(v,4) for k,v in d["kernel_ms_per_step"].items() if isinstance(v,float)}, round(d["e2e"]["value"]/1e9,1))
PY
done
