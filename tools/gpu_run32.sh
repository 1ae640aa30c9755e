set -x
mkdir -p gpurun_out
timeout 300 python tools/bench_ac.py HalfCheetah-v5 8192 3 2 > gpurun_out/r01_bench_ac_hc_v2.json 2>/dev/null; cat gpurun_out/r01_bench_ac_hc_v2.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r01_launches_ac.csv python tools/bench_ac.py HalfCheetah-v5 8192 1 1 > gpurun_out/ncu_ac.log 2>&1
tail -2 gpurun_out/ncu_ac.log | cut -c1-200
