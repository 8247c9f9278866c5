set -x
mkdir -p gpurun_out/r2
timeout 900 python -m pytest tests/test_gpu_forest.py tests/test_gpu_configs.py tests/test_gpu_subtract.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2/gputest_forest.txt
cat gpurun_out/r2/gputest_forest.txt
C3_ROWS=10000000 C3_TREES=16 timeout 600 python tools/config_probe.py c3 > gpurun_out/r2/c3_probe_v1.txt 2>&1
cat gpurun_out/r2/c3_probe_v1.txt
C3_ROWS=10000000 C3_TREES=16 timeout 900 ncu --set full --import-source on --clock-control none -k regex:hist_forest -c 6 -o gpurun_out/r2/c3_forest_v1 python tools/config_probe.py c3 > gpurun_out/r2/c3_ncu_v1.log 2>&1
ls -la gpurun_out/r2
