#!/bin/bash
# round 2, call 43 (1 GPU): tile-ahead prefetch distance of hist_rot_kernel in its final form (AVT_ROT_PF), C3 at 20 M rows
O=gpurun_out/r2; mkdir -p $O
for pf in 4 8 16 32 4; do
  AVT_ROT_PF=$pf timeout 600 python bench.py --rows 20000000 --steps 3 --warmup 3 --no-verify --no-cpu-baseline > $O/pf2_$pf.json 2> $O/pf2_$pf.err
  python - $O/pf2_$pf.json $pf <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print('pf',sys.argv[2],'ms_per_step',round(d['ms_per_step'],2),'hist',round(d['kernel_ms_per_step']['histogram'],2))
P
done
