# 2-GPU box: peer-memory exchange parity test, then bench A/B (peer path vs ncclAllReduce)
set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r01_topo_g2.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_peer_exchange.py -x -q > gpurun_out/r01_pytest_peer.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r01_pytest_peer.log
for mode in "PB_PEER_AR=1" "PB_PEER_AR=0"; do
  env $mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r01_bench_g2_$mode.out 2> gpurun_out/r01_bench_g2_$mode.err
  echo "rc=$? $mode"; grep -h "gradient exchange" gpurun_out/r01_bench_g2_$mode.err
  grep -h '^{' gpurun_out/r01_bench_g2_$mode.out | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['n_gpus'], d['phase_ms'], d['gpu_launches'])" || tail -5 gpurun_out/r01_bench_g2_$mode.err
done
env PB_UPDATE_TIMING=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 2 --warmup 3 --no-cpu-baseline 2>&1 >/dev/null | grep "update timing" | tail -4 | tee gpurun_out/r01_update_timing_g2_peer.txt
env PB_PEER_AR=0 PB_UPDATE_TIMING=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 2 --warmup 3 --no-cpu-baseline 2>&1 >/dev/null | grep "update timing" | tail -4 | tee gpurun_out/r01_update_timing_g2_nccl.txt
