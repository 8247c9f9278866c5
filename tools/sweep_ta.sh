for t in 4096 8192 16384 32768; do
echo "TA=$t"; AVT_HIST_TARGET_ADDRS=$t python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['kernel_ms_per_step']['histogram'], d['kernel_ms_per_step']['scan_plan'], d['roofline']['launch'])"
done
