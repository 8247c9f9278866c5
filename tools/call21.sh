#!/bin/bash
# round 2, call 21 (8 GPUs): C3 strong scaling at 8 and 4 GPUs as the driver launches it, 2-GPU tests
O=gpurun_out/r2; mkdir -p $O
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n --steps 5 --warmup 3 > $O/bench_c3_n$n.out 2> $O/bench_c3_n$n.err
  echo "rc n=$n $?"; grep '^{' $O/bench_c3_n$n.out > $O/bench_c3_n$n.json
  python - $O/bench_c3_n$n.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print(d['n_gpus'], 'value', round(d['value']/1e9,1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']/1e9,1), d['kernel_ms_per_step'], d['verify'])
P
  tail -3 $O/bench_c3_n$n.err
done
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -2
