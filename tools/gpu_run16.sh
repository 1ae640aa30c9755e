set -x
for mode in "PB_WG_OVERLAP=1" "PB_WG_OVERLAP=0" "PB_DW1_TC=0"; do
  echo "== $mode"
  env $mode PB_UPDATE_TIMING=1 timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 >/dev/null | grep "update timing" | tail -2
done > gpurun_out/r01_update_timing.txt 2>&1
cat gpurun_out/r01_update_timing.txt
