#!/bin/bash
# round 2, call 20 (1 GPU): launch list of the C5 step (which kernel runs at which level), C1 line with the tool legs
O=gpurun_out/r2; mkdir -p $O
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c5.csv python bench.py --workload c5 --rows 5000000 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_c5.log 2>&1; echo "rc $?"
timeout 900 python bench.py --workload c1 > $O/bench_c1_b.json 2> $O/bench_c1_b.err; echo "rc c1 $?"; tail -c 300 $O/bench_c1_b.err
python - <<'P'
import json
d=json.load(open('gpurun_out/r2/bench_c1_b.json')); print(json.dumps(d.get('e2e_tool'))); print(json.dumps(d.get('e2e')))
P
