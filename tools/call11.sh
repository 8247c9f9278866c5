set -x
mkdir -p gpurun_out/r2
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -10 > gpurun_out/r2/gputest_all_c.txt
cat gpurun_out/r2/gputest_all_c.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'hist_|scan_|repartition|inherit' -c 400 --csv --log-file gpurun_out/r2/launches_c3_10m_rot3.csv python bench.py --rows 10000000 --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2/ncu_c3_rot3.log 2>&1
timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/r2/bench_c3_full_rot3.json 2> gpurun_out/r2/bench_c3_full_rot3.err
tail -c 1500 gpurun_out/r2/bench_c3_full_rot3.json; tail -5 gpurun_out/r2/bench_c3_full_rot3.err
timeout 600 python bench.py --workload c2 --steps 10 --warmup 3 > gpurun_out/r2/bench_c2_joint3.json 2> gpurun_out/r2/bench_c2_joint3.err
tail -c 1500 gpurun_out/r2/bench_c2_joint3.json; tail -5 gpurun_out/r2/bench_c2_joint3.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2/launches_c2_joint3.csv python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r2/ncu_c2c.log 2>&1
