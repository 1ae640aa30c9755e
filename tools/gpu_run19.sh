# 1 GPU: wgrad_small_kernel (cp.async ring, co-resident with dw1_tc_kernel) parity + timing A/B
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "ppo or golden or cli or trainer" > gpurun_out/r01_pytest_wgs.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r01_pytest_wgs.log
{
for mode in "PB_WG_SMALL=1 PB_WG_OVERLAP=1" "PB_WG_SMALL=1 PB_WG_OVERLAP=0" "PB_WG_SMALL=0 PB_WG_OVERLAP=1"; do
  echo "== $mode"
  env $mode PB_UPDATE_TIMING=1 timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 >/dev/null | grep "update timing" | tail -1
  env $mode timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['phase_ms'], d['roofline_fp32']['frac'])"
done
} > gpurun_out/r01_update_wgs.txt 2>&1
cat gpurun_out/r01_update_wgs.txt
