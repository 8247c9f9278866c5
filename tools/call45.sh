#!/bin/bash
# round 2, call 45 (1 GPU): flat plan_level -- full GPU suite, C3 at 20 M rows (scan + plan time), C5
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --rows 20000000 --steps 3 --warmup 3 --no-cpu-baseline > $O/plan_c3.json 2> $O/plan_c3.err
python - $O/plan_c3.json <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print('c3 20M ms_per_step',round(d['ms_per_step'],2),d['kernel_ms_per_step'], d['verify']['result'][:30])
P
timeout 600 python bench.py --workload c5 --steps 3 --no-cpu-baseline > $O/plan_c5.json 2> $O/plan_c5.err
python - $O/plan_c5.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print('c5 ms_per_step',round(d['ms_per_step'],2),d['kernel_ms_per_step'], d['verify']['result'][:30])
P
