#!/bin/bash
# round 2, call 30 (1 GPU): what the driver runs at round end -- full GPU suite, smoke, bench.py as launched by the driver, the reference arm
O=gpurun_out/r2; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/gputest_final.txt 2>&1; tail -3 $O/gputest_final.txt
python -c "import __graft_entry__ as g; g.build(); g.smoke()" 2>&1 | tail -2
/usr/bin/time -v python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_final_n1.json 2> $O/bench_final_n1.err; echo "rc $?"; grep -E "Elapsed|Maximum resident" $O/bench_final_n1.err
python - $O/bench_final_n1.json <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print('c3 ms_per_step',round(d['ms_per_step'],1),'value',round(d['value']/1e9,1),'e2e',round(d['e2e']['value']/1e9,1),d['roofline']['frac'],d['roofline']['shared_atomics']['frac'],d['clocks'])
P
python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 > $O/bench_final_ref.json 2> $O/bench_final_ref.err; echo "rc ref $?"; head -c 300 $O/bench_final_ref.json; echo
