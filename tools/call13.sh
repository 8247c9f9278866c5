set -x
mkdir -p gpurun_out/r2
timeout 1200 python -m pytest tests/test_gpu_multi.py tests/test_gpu_kernels.py -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2/gputest_multi2.txt
cat gpurun_out/r2/gputest_multi2.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --rows 20000000 --trees 40 --steps 3 --warmup 3 > gpurun_out/r2/bench_c3_2gpu_small2.json 2> gpurun_out/r2/bench_c3_2gpu_small2.err
tail -c 900 gpurun_out/r2/bench_c3_2gpu_small2.json; tail -3 gpurun_out/r2/bench_c3_2gpu_small2.err
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2/bench_c3_2gpu_full.json 2> gpurun_out/r2/bench_c3_2gpu_full.err
tail -c 900 gpurun_out/r2/bench_c3_2gpu_full.json; tail -3 gpurun_out/r2/bench_c3_2gpu_full.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > gpurun_out/r2/bench_c3_ref_arm.json 2> gpurun_out/r2/bench_c3_ref_arm.err
tail -c 700 gpurun_out/r2/bench_c3_ref_arm.json
