set -x
mkdir -p gpurun_out/r2
timeout 1200 python -m pytest tests/test_gpu_forest.py tests/test_gpu_strategy.py tests/test_gpu_configs.py tests/test_gpu_subtract.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2/gputest_rot.txt
cat gpurun_out/r2/gputest_rot.txt
AVT_FOREST_ROT=0 timeout 600 python -m pytest tests/test_gpu_forest.py -m gpu -x -q 2>&1 | tail -3
C3_ROWS=10000000 C3_TREES=16 timeout 600 python tools/config_probe.py c3 > gpurun_out/r2/c3_probe_rot.txt 2>&1
cat gpurun_out/r2/c3_probe_rot.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'hist_|scan_|repartition|inherit' -c 400 --csv --log-file gpurun_out/r2/launches_c3_10m_rot.csv python bench.py --rows 10000000 --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2/ncu_c3_rot.log 2>&1
timeout 1500 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2/bench_c3_full_rot.json 2> gpurun_out/r2/bench_c3_full_rot.err
tail -c 1800 gpurun_out/r2/bench_c3_full_rot.json; tail -5 gpurun_out/r2/bench_c3_full_rot.err
C3_ROWS=10000000 C3_TREES=16 timeout 900 ncu --set full --import-source on --clock-control none -k regex:hist_rot -c 5 -o gpurun_out/r2/c3_rot_full python tools/config_probe.py c3 > gpurun_out/r2/c3_ncu_rot.log 2>&1
