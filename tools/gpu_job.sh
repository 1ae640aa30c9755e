mkdir -p gpurun_out/curves
# 1. final-kernel ncu capture at the pinned state + launch list of one bench iteration
M=smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_active.avg
timeout 600 ncu --set full --metrics $M --clock-control none --import-source on -k regex:env_step_kernel --launch-skip 264 -c 1 -o gpurun_out/r02_envstep_walk_final python tools/profile_walking.py Humanoid-v4 8192 tests/golden/walking_Humanoid-v4.npz 8 1 > gpurun_out/ncu_walk_final.log 2>&1
tail -1 gpurun_out/ncu_walk_final.log | cut -c1-120
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1200 -c 520 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
tail -1 gpurun_out/ncu_launches.log | cut -c1-120
# 2. compute-sanitizer
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python tools/sanitize_env.py 6 trainer > gpurun_out/r02_sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?"
timeout 1200 compute-sanitizer --tool racecheck --error-exitcode 7 python tools/sanitize_env.py 2 > gpurun_out/r02_sanitizer_racecheck.log 2>&1; echo "racecheck rc=$?"
tail -4 gpurun_out/r02_sanitizer_memcheck.log; tail -4 gpurun_out/r02_sanitizer_racecheck.log
# 3. Humanoid-v4 learning curves, 5 seeds to 1 M steps: 64 envs x 32 steps (batch 2048 and minibatch 64 as the reference's 1 x 2048)
for seed in 1 2 3 4 5; do
  ./build/ppo_continuous_action --env_id Humanoid-v4 --num_envs 64 --num_steps 32 --seed $seed --exp_name_stem curve_hum64 --num_eval_runs 10 > gpurun_out/curves/hum64_$seed.log 2>&1 &
done
wait
for seed in 1 2 3 4 5; do cp models/curve_hum64_$seed/scalars.jsonl gpurun_out/curves/curve_hum64_$seed.jsonl; tail -2 gpurun_out/curves/hum64_$seed.log | cut -c1-150; done
