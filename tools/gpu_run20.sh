# 1 GPU: CUDA-graph update A/B, parity, ncu of wgrad_small_kernel
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "ppo or golden or cli or trainer" > gpurun_out/r01_pytest_graph.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01_pytest_graph.log
{
for mode in "PB_UPDATE_GRAPH=1" "PB_UPDATE_GRAPH=0" "PB_UPDATE_GRAPH=1" "PB_UPDATE_GRAPH=0"; do
  echo "== $mode"
  env $mode timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['phase_ms'], round(d['e2e']['value']))"
done
} > gpurun_out/r01_update_graph.txt 2>&1
cat gpurun_out/r01_update_graph.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wgrad_small --launch-skip 400 -c 1 -o gpurun_out/r01_wgrad_small python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_wgs.log 2>&1
tail -2 gpurun_out/ncu_wgs.log | cut -c1-200
