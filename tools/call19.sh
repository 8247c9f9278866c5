#!/bin/bash
# round 2, call 19 (1 GPU): where the dead pairs' updates land (probe), phantom slot spread over classes (variant), tool tests with the device encoder
O=gpurun_out/r2; mkdir -p $O
cp avenir_b200/libavenir_tree.so /tmp/main.so
python tools/rot_diag.py 2>&1 | tail -4
run() { timeout 600 python bench.py --rows 20000000 --steps 3 --warmup 3 --no-verify --no-cpu-baseline > $O/ph_$1.json 2> $O/ph_$1.err
  python - $O/ph_$1.json $1 <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(sys.argv[2],'ms_per_step',round(d['ms_per_step'],2),'hist',round(d['kernel_ms_per_step']['histogram'],2))
P
}
run main
cp tools/variant_phcls.so avenir_b200/libavenir_tree.so; touch avenir_b200/libavenir_tree.so
python tools/rot_diag.py 2>&1 | tail -4
run phcls
cp /tmp/main.so avenir_b200/libavenir_tree.so; touch avenir_b200/libavenir_tree.so
run main2
python -m pytest tests/test_gpu_tool.py tests/test_host_vs_literal.py tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
