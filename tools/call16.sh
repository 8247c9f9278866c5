#!/bin/bash
# round 2, call 16 (1 GPU): hist_rot_kernel threads per block A/B (512 / 768 / 1024) on C3 at 20 M rows x 100 trees
O=gpurun_out/r2; mkdir -p $O
for th in 512 768 1024 512 768 1024; do
  AVT_ROT_THREADS=$th timeout 600 python bench.py --rows 20000000 --steps 3 --warmup 3 --no-verify --no-cpu-baseline > $O/th_$th.json 2> $O/th_$th.err
  python - $O/th_$th.json $th <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print('threads',sys.argv[2],'ms_per_step',round(d['ms_per_step'],2),'hist',round(d['kernel_ms_per_step']['histogram'],2))
P
done
for th in 768 1024; do AVT_ROT_THREADS=$th python -m pytest tests/test_gpu_forest.py tests/test_gpu_configs.py tests/test_gpu_strategy.py -m gpu -x -q 2>&1 | tail -2; done
