#!/bin/bash
# round 2, call 23 (1 GPU): batched loads in hist_sorted_kernel, repartition_rot_kernel; parity tests, C5 and C3 lines
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_configs.py tests/test_gpu_forest.py tests/test_gpu_strategy.py tests/test_gpu_subtract.py -m gpu -x -q 2>&1 | tail -4
timeout 900 python bench.py --workload c5 --steps 3 > $O/bench_c5_sorted.json 2> $O/bench_c5_sorted.err; echo "rc $?"; tail -c 300 $O/bench_c5_sorted.err
python - $O/bench_c5_sorted.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print('c5 ms_per_step',round(d['ms_per_step'],1),d['kernel_ms_per_step'],d['roofline']['frac'],d['e2e'])
P
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_c3_n1_b.json 2> $O/bench_c3_n1_b.err; echo "rc c3 $?"
python - $O/bench_c3_n1_b.json <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print('c3 ms_per_step',round(d['ms_per_step'],1),d['kernel_ms_per_step'],d['roofline']['frac'],d['e2e'],d['verify'])
P
