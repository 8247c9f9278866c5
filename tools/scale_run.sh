#!/bin/bash
# 1/2/4/8-GPU weak-scaling runs of bench.py on one box (peer exchange; NCCL variant at 8 for comparison)
mkdir -p gpurun_out
python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale_1.json
for n in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale_$n.json
done
AVT_BENCH_NCCL=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29650 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale_8_nccl.json
for f in 1 2 4 8 8_nccl; do python - <<PY
import json
d=json.load(open("gpurun_out/scale_$f.json"))
print("$f", round(d["value"]/1e9,1), round(d["ms_per_step"],4), {k:round(v,4) for k,v in d["kernel_ms_per_step"].items() if isinstance(v,float)}, round(d["e2e"]["value"]/1e9,1))
PY
done
