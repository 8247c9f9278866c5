set -x
mkdir -p gpurun_out/r2
nvidia-smi -L
timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2/gputest_multi.txt
cat gpurun_out/r2/gputest_multi.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --rows 20000000 --trees 40 --steps 3 --warmup 3 > gpurun_out/r2/bench_c3_2gpu_small.json 2> gpurun_out/r2/bench_c3_2gpu_small.err
tail -c 1500 gpurun_out/r2/bench_c3_2gpu_small.json; tail -8 gpurun_out/r2/bench_c3_2gpu_small.err
timeout 900 python bench.py --gpus 1 --rows 20000000 --trees 40 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2/bench_c3_1gpu_small.json 2> gpurun_out/r2/bench_c3_1gpu_small.err
tail -c 600 gpurun_out/r2/bench_c3_1gpu_small.json
timeout 600 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2/bench_c2_ab_joint.json 2>/dev/null
AVT_HIST_JOINT=0 timeout 600 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2/bench_c2_ab_old.json 2>/dev/null
python - <<'PY'
import json
for f in ['bench_c2_ab_joint','bench_c2_ab_old']:
    d=json.loads(open(f'gpurun_out/r2/{f}.json').read().strip().splitlines()[-1])
    print(f, d['ms_per_step'], d['kernel_ms_per_step']['histogram'], d['roofline']['kernel'])
PY
