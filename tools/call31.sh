#!/bin/bash
# round 2, call 31 (8 GPUs): final C3 strong-scaling set with the final build, launched as the driver launches it; 2-GPU parity tests
O=gpurun_out/r2; mkdir -p $O
python bench.py --gpus 1 --steps 20 --warmup 5 > $O/final_c3_n1.out 2> $O/final_c3_n1.err; echo "rc n=1 $?"
for n in 2 4 8; do
  timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800+n)) bench.py --gpus $n --steps 20 --warmup 5 > $O/final_c3_n$n.out 2> $O/final_c3_n$n.err; echo "rc n=$n $?"
done
for n in 1 2 4 8; do
  grep '^{' $O/final_c3_n$n.out > $O/final_c3_n$n.json
  python - $O/final_c3_n$n.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print(d['n_gpus'], 'value', round(d['value']/1e9,1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']/1e9,1), 'frac', round(d['roofline']['frac'],4), d['verify']['result'][:40], d['clocks']['reasons'])
P
done
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -v 2>&1 | tail -9 > $O/gputest_multi_final.txt; cat $O/gputest_multi_final.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29899 bench.py --gpus 8 --impl reference --steps 2 --warmup 1 > $O/final_ref_n8.out 2>/dev/null; grep '^{' $O/final_ref_n8.out | head -c 300; echo
