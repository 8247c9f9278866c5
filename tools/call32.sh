#!/bin/bash
# round 2, call 32 (1 GPU): ncu --set full of the final hist_rot_kernel (C3, 10 M rows) and hist_sorted_kernel (C5, 5 M rows), after plain runs of the same commands
O=gpurun_out/r2; mkdir -p $O
python bench.py --rows 10000000 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/plain_c3_10m.json 2>/dev/null; echo "rc plain c3 $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hist_rot --launch-skip 18 -c 6 -o $O/c3_rot_final2 python bench.py --rows 10000000 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_full_c3b.log 2>&1; echo "rc ncu c3 $?"
python bench.py --workload c5 --rows 5000000 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/plain_c5_5m.json 2>/dev/null; echo "rc plain c5 $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hist_sorted --launch-skip 21 -c 7 -o $O/c5_sorted_final python bench.py --workload c5 --rows 5000000 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_full_c5.log 2>&1; echo "rc ncu c5 $?"
ls -la $O/*.ncu-rep | tail -3
