#!/bin/bash
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_c3_n1_c.json 2> $O/bench_c3_n1_c.err; echo "rc c3 $?"
python - $O/bench_c3_n1_c.json <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print('c3 ms_per_step',round(d['ms_per_step'],1),d['kernel_ms_per_step'],d['roofline']['frac'],round(d['e2e']['value']/1e9,1))
P
