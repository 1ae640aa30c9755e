set -x
N=${N:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/r01_bench_g$N.json 2> gpurun_out/r01_bench_g$N.err
cat gpurun_out/r01_bench_g$N.json; tail -5 gpurun_out/r01_bench_g$N.err
