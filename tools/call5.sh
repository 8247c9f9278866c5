set -x
mkdir -p gpurun_out/r2
timeout 1200 python -m pytest tests/test_gpu_forest.py tests/test_gpu_strategy.py tests/test_gpu_configs.py tests/test_gpu_subtract.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -12 > gpurun_out/r2/gputest_v5.txt
cat gpurun_out/r2/gputest_v5.txt
C3_ROWS=10000000 C3_TREES=16 timeout 600 python tools/config_probe.py c3 > gpurun_out/r2/c3_probe_v5.txt 2>&1
cat gpurun_out/r2/c3_probe_v5.txt
timeout 1500 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2/bench_c3_full_v5.json 2> gpurun_out/r2/bench_c3_full_v5.err
tail -c 1800 gpurun_out/r2/bench_c3_full_v5.json; tail -5 gpurun_out/r2/bench_c3_full_v5.err
C3_ROWS=10000000 C3_TREES=16 timeout 900 ncu --set full --import-source on --clock-control none -k regex:hist_forest -c 6 -o gpurun_out/r2/c3_forest_v5 python tools/config_probe.py c3 > gpurun_out/r2/c3_ncu_v5.log 2>&1
