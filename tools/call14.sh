#!/bin/bash
# round 2, call 14 (1 GPU): full GPU suite; secondary bench lines (c5, c4, c1); final C3 line; DRAM traffic + launch list + full ncu of hist_rot
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/gputest_all_d.txt 2>&1; tail -3 $O/gputest_all_d.txt
for w in c1 c4 c5; do
  timeout 900 python bench.py --workload $w > $O/bench_$w.json 2> $O/bench_$w.err; echo "rc $w $?"; tail -c 400 $O/bench_$w.err; head -c 600 $O/bench_$w.json; echo
done
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_c3_n1_final.json 2> $O/bench_c3_n1_final.err; echo "rc c3 $?"; head -c 1500 $O/bench_c3_n1_final.json; echo
# DRAM bytes per launch of hist_rot on the bench configuration (one pass of two counters; 18 warm-up launches skipped)
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:hist_rot --launch-skip 18 -c 6 --csv \
  --log-file $O/traffic_c3_full.csv python bench.py --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_traffic_c3.log 2>&1; echo "rc traffic $?"
# launch list of the same command
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3_full.csv \
  python bench.py --steps 2 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_launches_c3.log 2>&1; echo "rc launches $?"
# full capture at 10 M rows x 100 trees (ncu restores device memory between its ~40 passes: the 100 M-row state is too large for that)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hist_rot --launch-skip 18 -c 6 -o $O/c3_rot_final \
  python bench.py --rows 10000000 --steps 1 --warmup 3 --no-verify --no-cpu-baseline > $O/ncu_full_c3.log 2>&1; echo "rc full $?"
ls -la $O | tail -12
