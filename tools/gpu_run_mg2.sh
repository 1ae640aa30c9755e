set -x
for mode in "PB_DW1_TC=1" "PB_DW1_TC=0"; do
  env $mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_g2_$mode.out 2> gpurun_out/r01_bench_g2_$mode.err
  echo "rc=$? $mode"; grep -h '^{' gpurun_out/r01_bench_g2_$mode.out | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['n_gpus'], d['phase_ms'], d['gpu_launches'])" || tail -5 gpurun_out/r01_bench_g2_$mode.err
done
