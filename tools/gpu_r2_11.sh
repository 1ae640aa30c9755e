mkdir -p gpurun_out
M=smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_active.avg
timeout 600 ncu --set full --metrics $M --clock-control none --import-source on -k regex:env_step_kernel --launch-skip 264 -c 1 -o gpurun_out/r02_envstep_walk_v4 python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 8 1 > gpurun_out/ncu_walk_v4.log 2>&1
tail -2 gpurun_out/ncu_walk_v4.log | cut -c1-200
