#!/bin/bash
# round 2, call 26 (1 GPU): live-pair compaction (AVT_ROT_RING) -- parity tests, then C3 at 20 M rows with the rings off / fused levels / all levels
O=gpurun_out/r2; mkdir -p $O
for m in 1 2; do AVT_ROT_RING=$m timeout 900 python -m pytest tests/test_gpu_forest.py tests/test_gpu_configs.py tests/test_gpu_strategy.py tests/test_gpu_subtract.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -3; done
for m in 0 1 2 0 1 2; do
  AVT_ROT_RING=$m timeout 600 python bench.py --rows 20000000 --steps 3 --warmup 3 --no-cpu-baseline > $O/ring_$m.json 2> $O/ring_$m.err
  python - $O/ring_$m.json $m <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print('ring',sys.argv[2],'ms_per_step',round(d['ms_per_step'],2),'hist',round(d['kernel_ms_per_step']['histogram'],2), d['verify']['result'][:30])
P
done
