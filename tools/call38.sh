#!/bin/bash
# round 2, call 38 (8 GPUs): the whole GPU suite on a multi-GPU box (no skips), then the final C3 strong-scaling set with the last build
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests -m gpu -q > $O/gputest_final8.txt 2>&1; tail -3 $O/gputest_final8.txt
python bench.py --gpus 1 --steps 20 --warmup 5 > $O/final2_c3_n1.out 2> $O/final2_c3_n1.err; echo "rc n=1 $?"
for n in 2 4 8; do
  timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29900+n)) bench.py --gpus $n --steps 20 --warmup 5 > $O/final2_c3_n$n.out 2> $O/final2_c3_n$n.err; echo "rc n=$n $?"
done
for n in 1 2 4 8; do
  grep '^{' $O/final2_c3_n$n.out > $O/final2_c3_n$n.json
  python - $O/final2_c3_n$n.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print(d['n_gpus'], 'value', round(d['value']/1e9,1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']/1e9,1), 'frac', round(d['roofline']['frac'],4), d['verify']['result'][:40], d['clocks']['reasons'], d['kernel_ms_per_step']['scan_plan'])
P
done
