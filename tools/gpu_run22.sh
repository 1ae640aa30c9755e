# 2-GPU box: fused slab-reduction + exchange kernel: parity (fused and two-kernel variants), bench A/B
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_peer_exchange.py -x -q > gpurun_out/r01_pytest_peer_fused.log 2>&1; echo "pytest fused rc=$?"; tail -5 gpurun_out/r01_pytest_peer_fused.log
PB_PEER_FUSED=0 timeout 900 python -m pytest tests/test_gpu_peer_exchange.py -x -q > gpurun_out/r01_pytest_peer_2k.log 2>&1; echo "pytest two-kernel rc=$?"; tail -3 gpurun_out/r01_pytest_peer_2k.log
for mode in "PB_PEER_FUSED=1" "PB_PEER_FUSED=0" "PB_PEER_FUSED=1" "PB_PEER_FUSED=0"; do
  env $mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_g2c_$mode.out 2> gpurun_out/r01_bench_g2c_$mode.err
  echo "rc=$? $mode"; grep -h "gradient exchange" gpurun_out/r01_bench_g2c_$mode.err
  grep -h '^{' gpurun_out/r01_bench_g2c_$mode.out | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['n_gpus'], d['phase_ms'], d['gpu_launches'], d['gradient_exchange'])" || tail -5 gpurun_out/r01_bench_g2c_$mode.err
done
timeout 300 python tools/bench_ac.py > gpurun_out/r01_bench_ac_hc.json 2> gpurun_out/r01_bench_ac_hc.err; cat gpurun_out/r01_bench_ac_hc.json; tail -2 gpurun_out/r01_bench_ac_hc.err
