#!/bin/bash
# round 2, call 17 (1 GPU): DataPartitioner / encode tests on the GPU, full C3 line with the prefetch
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests/test_dap.py tests/test_encode.py tests/test_crc.py -m gpu -x -q 2>&1 | tail -4
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_c3_n1_pf.json 2> $O/bench_c3_n1_pf.err; echo "rc c3 $?"; head -c 400 $O/bench_c3_n1_pf.json; echo
