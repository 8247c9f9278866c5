set -x
mkdir -p gpurun_out/r2
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/r2/gputest_all_a.txt
cat gpurun_out/r2/gputest_all_a.txt
C3_ROWS=10000000 C3_TREES=16 timeout 600 python tools/config_probe.py c3 > gpurun_out/r2/c3_probe_v6.txt 2>&1
cat gpurun_out/r2/c3_probe_v6.txt
timeout 600 python bench.py --workload c2 --steps 10 --warmup 3 > gpurun_out/r2/bench_c2_joint.json 2> gpurun_out/r2/bench_c2_joint.err
tail -c 2500 gpurun_out/r2/bench_c2_joint.json; tail -5 gpurun_out/r2/bench_c2_joint.err
AVT_HIST_VARIANT_OLD=1 timeout 600 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2/bench_c2_old.json 2> gpurun_out/r2/bench_c2_old.err
timeout 1500 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2/bench_c3_full_v6.json 2> gpurun_out/r2/bench_c3_full_v6.err
tail -c 1800 gpurun_out/r2/bench_c3_full_v6.json; tail -5 gpurun_out/r2/bench_c3_full_v6.err
