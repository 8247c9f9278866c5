#!/bin/bash
# round 2, call 22 (1 GPU): hist_sorted_kernel parity tests, C5 with the sorted-rows kernel off / fallback-only / whenever rot does not fit
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_configs.py tests/test_gpu_subtract.py -m gpu -x -q 2>&1 | tail -4
for m in 0 1 2; do
  AVT_HIST_SORTED=$m timeout 900 python bench.py --workload c5 --steps 3 --no-cpu-baseline > $O/c5_sorted_$m.json 2> $O/c5_sorted_$m.err; echo "rc $?"; tail -c 300 $O/c5_sorted_$m.err
  python - $O/c5_sorted_$m.json $m <<'P'
import json,sys
d=json.load(open(sys.argv[1]))
print('sorted mode',sys.argv[2],'ms_per_step',round(d['ms_per_step'],1),d['kernel_ms_per_step'],d['roofline']['kernel'],d['verify'])
P
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"hist_|scan_kernel|repartition|DeviceRadixSort|sorted_gather" -c 400 --csv --log-file $O/launches_c5_sorted.csv python bench.py --workload c5 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_c5_sorted.log 2>&1; echo "rc ncu $?"
