# 1 GPU: ncu --set full of the remaining update / rollout kernels (one launch each, bench state)
set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fwd_bwd_kernel|l1_forward_tc_kernel|policy_forward_kernel|grad_reduce_kernel|adam2_kernel" --launch-skip 1500 -c 6 -o gpurun_out/r01_update_v8 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_update_v8.log 2>&1
tail -3 gpurun_out/ncu_update_v8.log | cut -c1-200
