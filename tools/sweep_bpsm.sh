for b in 1 2 3 4; do for t in 1024 4096; do
echo "BPSM=$b TA=$t"; AVT_HIST_BPSM=$b AVT_HIST_TARGET_ADDRS=$t python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['kernel_ms_per_step']['histogram'], d['kernel_ms_per_step']['scan_plan'], d['roofline']['launch'])"
done; done
