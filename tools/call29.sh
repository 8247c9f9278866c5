#!/bin/bash
# round 2, call 29 (1 GPU): C2 with the joint-code kernel only on small frontiers (hybrid), same box A/B
O=gpurun_out/r2; mkdir -p $O
run() { timeout 600 env $1 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline > $O/c2h_$2.json 2> $O/c2h_$2.err
  python - $O/c2h_$2.json "$2" <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(sys.argv[2],'ms_per_step',round(d['ms_per_step'],4),d['kernel_ms_per_step'])
P
}
run "AVT_HIST_JOINT=0" off
run "AVT_HIST_JOINT=1 AVT_JOINT_MAX_NODES=4" j4
run "AVT_HIST_JOINT=1 AVT_JOINT_MAX_NODES=8" j8
run "AVT_HIST_JOINT=1 AVT_JOINT_MAX_NODES=16" j16
run "AVT_HIST_JOINT=1 AVT_JOINT_MAX_NODES=32" j32
run "AVT_HIST_JOINT=1" jall
run "AVT_HIST_JOINT=0" off2
AVT_HIST_JOINT=1 AVT_JOINT_MAX_NODES=8 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -2
