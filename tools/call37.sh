#!/bin/bash
# round 2, call 37 (1 GPU): randomised differential run of the level loop against the oracle (tools/fuzz_gpu.py), three seeds
O=gpurun_out/r2; mkdir -p $O
for seed in 1 2 3; do timeout 400 python tools/fuzz_gpu.py 100 $seed > $O/fuzz_$seed.txt 2>&1; echo "seed $seed rc $?"; tail -4 $O/fuzz_$seed.txt | cut -c1-400; done
