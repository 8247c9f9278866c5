set -x
mkdir -p gpurun_out/r2
timeout 900 python -m pytest tests/test_bagging.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2/gputest_new.txt
cat gpurun_out/r2/gputest_new.txt
timeout 900 python bench.py --rows 10000000 --trees 16 --steps 3 --warmup 3 > gpurun_out/r2/bench_c3_small.json 2> gpurun_out/r2/bench_c3_small.err
tail -c 3000 gpurun_out/r2/bench_c3_small.json; tail -5 gpurun_out/r2/bench_c3_small.err
timeout 1500 python bench.py --steps 3 --warmup 3 > gpurun_out/r2/bench_c3_full_base.json 2> gpurun_out/r2/bench_c3_full_base.err
tail -c 3000 gpurun_out/r2/bench_c3_full_base.json; tail -5 gpurun_out/r2/bench_c3_full_base.err
