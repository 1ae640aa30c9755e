# 8-GPU box: weak-scaling bench, peer-memory exchange (default) vs ncclAllReduce
set -x
mkdir -p gpurun_out
N=8
for mode in "PB_PEER_AR=1" "PB_PEER_AR=0"; do
  env $mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_g8_$mode.out 2> gpurun_out/r01_bench_g8_$mode.err
  echo "rc=$? $mode"; grep -h "gradient exchange" gpurun_out/r01_bench_g8_$mode.err
  grep -h '^{' gpurun_out/r01_bench_g8_$mode.out | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['n_gpus'], d['phase_ms'], d['gpu_launches'], d['gradient_exchange'], round(d['e2e']['value']))" || tail -8 gpurun_out/r01_bench_g8_$mode.err
done
