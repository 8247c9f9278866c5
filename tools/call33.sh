#!/bin/bash
# round 2, call 33 (1 GPU): hist_smem_kernel attribute-group size on C4 (wide categorical columns)
O=gpurun_out/r2; mkdir -p $O
for kb in 200 100 64 32 16 200; do
  AVT_HIST_BUDGET_KB=$kb timeout 600 python bench.py --workload c4 --no-cpu-baseline > $O/c4_b$kb.json 2> $O/c4_b$kb.err
  python - $O/c4_b$kb.json $kb <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print('budget KB',sys.argv[2],'ms_per_step',round(d['ms_per_step'],3),d['kernel_ms_per_step']['histogram'],d['roofline']['launch'])
P
done
