#!/bin/bash
# round 2, call 35 (1 GPU): A/B of hist_rot_kernel address arithmetic (old .so vs new .so), C3 at 20 M rows
O=gpurun_out/r2; mkdir -p $O
run() { cp tools/variant_$1.so avenir_b200/libavenir_tree.so; touch avenir_b200/libavenir_tree.so
  timeout 600 python bench.py --rows 20000000 --steps 3 --warmup 3 --no-cpu-baseline > $O/ab_$1.json 2> $O/ab_$1.err
  python - $O/ab_$1.json $1 <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(sys.argv[2],'ms_per_step',round(d['ms_per_step'],2),'hist',round(d['kernel_ms_per_step']['histogram'],2), d['verify']['result'][:30])
P
}
run old; run new; run old; run new
