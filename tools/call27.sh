#!/bin/bash
# round 2, call 27 (1 GPU): live-pair compaction with staged head codes
O=gpurun_out/r2; mkdir -p $O
AVT_ROT_RING=2 timeout 900 python -m pytest tests/test_gpu_forest.py tests/test_gpu_configs.py tests/test_gpu_strategy.py tests/test_gpu_subtract.py -m gpu -x -q 2>&1 | tail -3
for m in 0 1 2; do
  AVT_ROT_RING=$m timeout 600 python bench.py --rows 20000000 --steps 3 --warmup 3 --no-cpu-baseline > $O/ring2_$m.json 2> $O/ring2_$m.err
  python - $O/ring2_$m.json $m <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print('ring',sys.argv[2],'ms_per_step',round(d['ms_per_step'],2),'hist',round(d['kernel_ms_per_step']['histogram'],2), d['verify']['result'][:30])
P
done
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
