#!/bin/bash
# round 2, call 18 (1 GPU): hist_rot_kernel register lookahead depth A/B (2 / 4 / 6 / 8 trees) x L2 prefetch distance, C3 at 20 M rows
O=gpurun_out/r2; mkdir -p $O
cp avenir_b200/libavenir_tree.so /tmp/la4.so
run() { # $1 = label
  timeout 600 python bench.py --rows 20000000 --steps 3 --warmup 3 --no-verify --no-cpu-baseline > $O/la_$1.json 2> $O/la_$1.err
  python - $O/la_$1.json $1 <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(sys.argv[2],'ms_per_step',round(d['ms_per_step'],2),'hist',round(d['kernel_ms_per_step']['histogram'],2))
P
}
for la in 4 2 6 8; do
  if [ $la = 4 ]; then cp /tmp/la4.so avenir_b200/libavenir_tree.so; else cp tools/variants_la$la.so avenir_b200/libavenir_tree.so; fi
  touch avenir_b200/libavenir_tree.so
  for pf in 4 12; do AVT_ROT_PF=$pf run "la${la}_pf${pf}"; done
done
cp /tmp/la4.so avenir_b200/libavenir_tree.so; touch avenir_b200/libavenir_tree.so
python -m pytest tests/test_gpu_forest.py tests/test_gpu_configs.py tests/test_gpu_strategy.py -m gpu -x -q 2>&1 | tail -2
