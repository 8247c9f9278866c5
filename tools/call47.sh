#!/bin/bash
O=gpurun_out/r2; mkdir -p $O
python bench.py --gpus 1 --steps 20 --warmup 5 > $O/final3_c3_n1.out 2> $O/final3_c3_n1.err; echo "rc $?"
grep '^{' $O/final3_c3_n1.out > $O/final3_c3_n1.json
python - $O/final3_c3_n1.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print(d['n_gpus'], 'value', round(d['value']/1e9,1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']/1e9,1), d['kernel_ms_per_step'], d['verify']['result'][:40], d['roofline']['frac'])
P
