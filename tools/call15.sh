#!/bin/bash
# round 2, call 15 (1 GPU): hist_rot_kernel L2 prefetch distance A/B on C3 at 20 M rows x 100 trees
O=gpurun_out/r2; mkdir -p $O
for pf in 0 4 8 16 0 8; do
  AVT_ROT_PF=$pf timeout 600 python bench.py --rows 20000000 --steps 3 --warmup 3 --no-verify --no-cpu-baseline > $O/pf_$pf.json 2> $O/pf_$pf.err
  python - $O/pf_$pf.json $pf <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print('pf',sys.argv[2],'ms_per_step',round(d['ms_per_step'],2),'hist',round(d['kernel_ms_per_step']['histogram'],2))
P
done
AVT_ROT_PF=8 python -m pytest tests/test_gpu_forest.py tests/test_gpu_configs.py -m gpu -x -q 2>&1 | tail -2
