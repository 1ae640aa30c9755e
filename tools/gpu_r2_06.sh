set -x
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_v1.json 2> gpurun_out/r02_bench_v1.err
tail -3 gpurun_out/r02_bench_v1.err
cat gpurun_out/r02_bench_v1.json | cut -c1-3000
