#!/bin/bash
# round 2, call 24 (1 GPU): warp-per-split CPG sweep: parity tests and the C4 line
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_configs.py tests/test_gpu_tool.py tests/test_host_vs_literal.py tests/test_crc.py -m gpu -x -q 2>&1 | tail -4
timeout 900 python bench.py --workload c4 > $O/bench_c4_b.json 2> $O/bench_c4_b.err; echo "rc $?"; tail -c 300 $O/bench_c4_b.err
python - $O/bench_c4_b.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print('c4 ms_per_step',round(d['ms_per_step'],2),d['kernel_ms_per_step'],d['roofline']['frac'],d['e2e'],d['verify'])
P
