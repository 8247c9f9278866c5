#!/bin/bash
# round 2, call 44 (1 GPU): ncu --set full of the six scan_kernel launches of a C3 step (10 M rows x 100 trees)
O=gpurun_out/r2; mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:scan_kernel --launch-skip 18 -c 6 -o $O/c3_scan python bench.py --rows 10000000 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_scan.log 2>&1; echo "rc $?"
