set -x
{
for mode in "PB_WG_CARVEOUT=1 PB_WG_OVERLAP=1" "PB_WG_CARVEOUT=0 PB_WG_OVERLAP=1"; do
  echo "== $mode"
  env $mode PB_UPDATE_TIMING=1 timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 >/dev/null | grep "update timing" | tail -1
  env $mode timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), d['phase_ms'])"
done
} > gpurun_out/r01_update_carveout.txt 2>&1
cat gpurun_out/r01_update_carveout.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dw1_tc_kernel --launch-skip 100 -c 1 -o gpurun_out/r01_dw1_tc python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_dw1_full.log 2>&1
tail -3 gpurun_out/ncu_dw1_full.log | cut -c1-200
