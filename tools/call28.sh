#!/bin/bash
# round 2, call 28 (1 GPU): hist_sorted_kernel with staged row tiles: parity tests, C5 line, launch list
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_configs.py tests/test_gpu_subtract.py -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --workload c5 --steps 3 > $O/bench_c5_sorted2.json 2> $O/bench_c5_sorted2.err; echo "rc $?"; tail -c 300 $O/bench_c5_sorted2.err
python - $O/bench_c5_sorted2.json <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print('c5 ms_per_step',round(d['ms_per_step'],1),d['kernel_ms_per_step'],d['roofline']['frac'],d['roofline']['kernel'])
P
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"hist_|scan_kernel|repartition|DeviceRadixSort|sorted_gather" -c 400 --csv --log-file $O/launches_c5_sorted2.csv python bench.py --workload c5 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_c5_sorted2.log 2>&1; echo "rc ncu $?"
