#!/bin/bash
# round 2, call 42 (1 GPU): DRAM bytes of the histogram launches of the C5 and C4 steps (for roofline.traffic of those lines)
O=gpurun_out/r2; mkdir -p $O
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"hist_rot|hist_sorted" --launch-skip 36 -c 12 --csv --log-file $O/traffic_c5.csv python bench.py --workload c5 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_traffic_c5.log 2>&1; echo "rc c5 $?"
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:"hist_smem" -c 12 --csv --log-file $O/traffic_c4.csv python bench.py --workload c4 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_traffic_c4.log 2>&1; echo "rc c4 $?"
tail -3 $O/traffic_c5.csv | cut -c1-300; tail -3 $O/traffic_c4.csv | cut -c1-300
