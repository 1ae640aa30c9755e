set -x
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r01_launches_ac_v2.csv python tools/bench_ac.py HalfCheetah-v5 8192 1 1 > gpurun_out/ncu_ac.log 2>&1
tail -2 gpurun_out/ncu_ac.log | cut -c1-200
