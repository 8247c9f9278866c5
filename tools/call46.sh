#!/bin/bash
# round 2, call 46 (2 GPUs): flat plan_level under the two exchange forms -- 2-GPU parity tests and the C3 line at 2 GPUs
O=gpurun_out/r2; mkdir -p $O
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -2
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29951 bench.py --gpus 2 --steps 10 --warmup 3 > $O/plan_c3_n2.out 2> $O/plan_c3_n2.err; echo "rc $?"
grep '^{' $O/plan_c3_n2.out > $O/plan_c3_n2.json
python - $O/plan_c3_n2.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print(d['n_gpus'], 'value', round(d['value']/1e9,1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']/1e9,1), d['kernel_ms_per_step'], d['verify']['result'][:40])
P
