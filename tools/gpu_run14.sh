set -x
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r01_launches_dw1.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_dw1.log 2>&1
tail -2 gpurun_out/ncu_dw1.log | cut -c1-200
