#!/bin/bash
# round 2, call 34 (1 GPU): full GPU suite with the 100 KB attribute-group budget, C4 and C2 lines
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --workload c4 > $O/bench_c4_d.json 2> $O/bench_c4_d.err; echo "rc c4 $?"
timeout 600 python bench.py --workload c2 --steps 20 --warmup 5 > $O/bench_c2_d.json 2> $O/bench_c2_d.err; echo "rc c2 $?"
python - <<'P'
import json
d=json.load(open('gpurun_out/r2/bench_c4_d.json')); print('c4',round(d['ms_per_step'],3),d['kernel_ms_per_step'],round(d['roofline']['frac'],3))
d=json.loads([l for l in open('gpurun_out/r2/bench_c2_d.json') if l.startswith('{')][-1]); print('c2',round(d['ms_per_step'],4),round(d['roofline']['frac'],3),round(d['e2e']['value']/1e9,1))
P
