# 2-GPU box: peer exchange inside the CUDA-graph update (parity + bench), then 1-GPU A/B of the forward/backward tile height
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_peer_exchange.py -x -q > gpurun_out/r01_pytest_peer_graph.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r01_pytest_peer_graph.log
for mode in "PB_PEER_AR=1" "PB_PEER_AR=0"; do
  env $mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_g2b_$mode.out 2> gpurun_out/r01_bench_g2b_$mode.err
  echo "rc=$? $mode"; grep -h "gradient exchange" gpurun_out/r01_bench_g2b_$mode.err
  grep -h '^{' gpurun_out/r01_bench_g2b_$mode.out | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['n_gpus'], d['phase_ms'], d['gpu_launches'], d['gradient_exchange'])" || tail -5 gpurun_out/r01_bench_g2b_$mode.err
done
{
for mode in "PB_FB_TM=64" "PB_FB_TM=32" "PB_FB_TM=64" "PB_FB_TM=32"; do
  echo "== $mode"
  env $mode timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['phase_ms'], round(d['e2e']['value']))"
done
env PB_FB_TM=32 timeout 600 python -m pytest tests -m gpu -x -q -k "ppo or golden or trainer" 2>&1 | tail -2
} > gpurun_out/r01_update_tm.txt 2>&1
cat gpurun_out/r01_update_tm.txt
