#!/bin/bash
O=gpurun_out/r2; mkdir -p $O
timeout 900 python bench.py --workload c4 > $O/bench_c4_c.json 2> $O/bench_c4_c.err; echo "rc $?"; tail -c 300 $O/bench_c4_c.err
python - $O/bench_c4_c.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print('c4 ms_per_step',round(d['ms_per_step'],2),d['kernel_ms_per_step'],d['roofline']['frac'],d['e2e'],d['verify'])
P
