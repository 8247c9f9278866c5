set -x
mkdir -p gpurun_out/r2
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -10 > gpurun_out/r2/gputest_all_b.txt
cat gpurun_out/r2/gputest_all_b.txt
C3_ROWS=10000000 C3_TREES=16 timeout 600 python tools/config_probe.py c3 > gpurun_out/r2/c3_probe_rot2.txt 2>&1
cat gpurun_out/r2/c3_probe_rot2.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'hist_|scan_|repartition|inherit' -c 400 --csv --log-file gpurun_out/r2/launches_c3_10m_rot2.csv python bench.py --rows 10000000 --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2/ncu_c3_rot2.log 2>&1
timeout 1500 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2/bench_c3_full_rot2.json 2> gpurun_out/r2/bench_c3_full_rot2.err
tail -c 1500 gpurun_out/r2/bench_c3_full_rot2.json; tail -5 gpurun_out/r2/bench_c3_full_rot2.err
timeout 600 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2/bench_c2_joint2.json 2> gpurun_out/r2/bench_c2_joint2.err
tail -c 1500 gpurun_out/r2/bench_c2_joint2.json; tail -5 gpurun_out/r2/bench_c2_joint2.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2/launches_c2_joint2.csv python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r2/ncu_c2b.log 2>&1
