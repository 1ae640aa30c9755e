set -x
mkdir -p gpurun_out
python tools/make_walking_checkpoint.py gpurun_out 50 > gpurun_out/walking.log 2>&1
tail -12 gpurun_out/walking.log
python tools/profile_walking.py Humanoid-v4 8192 gpurun_out/walking_Humanoid-v4.npz 3 3 > gpurun_out/walk_time_hum.log 2>&1
python tools/profile_walking.py Ant-v5 16384 gpurun_out/walking_Ant-v5.npz 3 3 > gpurun_out/walk_time_ant.log 2>&1
cat gpurun_out/walk_time_hum.log gpurun_out/walk_time_ant.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
tail -3 gpurun_out/pytest_gpu.log
M=smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_active.avg
timeout 600 ncu --set full --metrics $M --clock-control none --import-source on -k regex:env_step_kernel --launch-skip 104 -c 1 -o gpurun_out/r02_envstep_walk_v0 python tools/profile_walking.py Humanoid-v4 8192 gpurun_out/walking_Humanoid-v4.npz 3 1 > gpurun_out/ncu_walk_v0.log 2>&1
tail -2 gpurun_out/ncu_walk_v0.log | cut -c1-300
ls -la gpurun_out
