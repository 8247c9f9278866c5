set -x
mkdir -p gpurun_out/r2
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
./tools/microbench/atoms_bench > gpurun_out/r2/atoms_bench.txt 2>&1
cat gpurun_out/r2/atoms_bench.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2/gputest_base.txt
cat gpurun_out/r2/gputest_base.txt
C3_ROWS=10000000 C3_TREES=16 timeout 600 python tools/config_probe.py c3 > gpurun_out/r2/c3_probe_base.txt 2>&1
cat gpurun_out/r2/c3_probe_base.txt
C5_ROWS=4000000 timeout 600 python tools/config_probe.py c5 > gpurun_out/r2/c5_probe_base.txt 2>&1
cat gpurun_out/r2/c5_probe_base.txt
C3_ROWS=10000000 C3_TREES=16 timeout 900 ncu --set full --clock-control none -k regex:hist_ -c 12 --csv --page raw --log-file gpurun_out/r2/c3_base_ncu_raw.csv python tools/config_probe.py c3 > gpurun_out/r2/c3_ncu.log 2>&1
C5_ROWS=4000000 timeout 900 ncu --set full --clock-control none -k regex:hist_ -c 8 --csv --page raw --log-file gpurun_out/r2/c5_base_ncu_raw.csv python tools/config_probe.py c5 > gpurun_out/r2/c5_ncu.log 2>&1
ls -la gpurun_out/r2
