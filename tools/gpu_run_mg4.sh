# 4-GPU box: weak-scaling bench at N=4 and N=2 with the final round-1 kernels
set -x
mkdir -p gpurun_out
for N in 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_g${N}_final.json 2> gpurun_out/r01_bench_g${N}_final.err
  echo "rc=$? N=$N"; grep -h '^{' gpurun_out/r01_bench_g${N}_final.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['n_gpus'], d['phase_ms'], d['gradient_exchange'], round(d['e2e']['value']))" || tail -8 gpurun_out/r01_bench_g${N}_final.err
done
