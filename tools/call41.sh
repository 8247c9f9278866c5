#!/bin/bash
# round 2, call 41 (1 GPU): last check of the final tree -- full GPU suite, smoke, default bench line, reference arm
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/gputest_last.txt 2>&1; tail -3 $O/gputest_last.txt
python -c "import __graft_entry__ as g; g.build(); g.smoke()" 2>&1 | tail -2
python bench.py > $O/bench_last.json 2> $O/bench_last.err; echo "rc bench $?"
python - <<'P'
import json
d=json.loads([l for l in open('gpurun_out/r2/bench_last.json') if l.startswith('{')][-1])
print(round(d['ms_per_step'],1), round(d['value']/1e9,1), round(d['e2e']['value']/1e9,1), d['roofline']['frac'], d['roofline']['kernel'], d['gpu_launches'], d['clocks'], d['cpu_baseline']['value'])
P
