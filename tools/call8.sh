set -x
mkdir -p gpurun_out/r2
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2/launches_c2_joint.csv python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r2/ncu_c2.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'hist_|scan_|repartition|inherit' -c 400 --csv --log-file gpurun_out/r2/launches_c3_10m.csv python bench.py --rows 10000000 --steps 1 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2/ncu_c3.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:hist_joint -s 8 -c 8 -o gpurun_out/r2/c2_joint_full python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r2/ncu_c2_full.log 2>&1
ls -la gpurun_out/r2 | tail -5
