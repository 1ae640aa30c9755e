set -x
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/r01_bench_v7.json 2> gpurun_out/r01_bench_v7.err; tail -2 gpurun_out/r01_bench_v7.err
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01_bench_ref_v7.json 2> gpurun_out/r01_bench_ref_v7.err; tail -2 gpurun_out/r01_bench_ref_v7.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r01_launches_v7.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log | cut -c1-200
