# Learning-curve evidence through the product CLI (profiles/r02_learning_curves_humanoid.txt): Humanoid-v4 to 1 M env steps, 5 seeds,
# all flags at the reference's defaults except the rollout shape (the reference's 1 env x 2048 steps is one warp here: ~400 env-steps/s).
# Compare with the reference's published TensorBoard screenshots (Humanoid-v4 ~600-650 at 1 M steps).
# gpurun --timeout 1500 -- 'bash tools/gpu_learning_curves.sh'
cd /root/repo
mkdir -p gpurun_out/curves
for shape in "8 256 curve_hum8" "64 32 curve_hum64"; do
  set -- $shape
  for seed in 1 2 3 4 5; do
    ./build/ppo_continuous_action --env_id Humanoid-v4 --num_envs $1 --num_steps $2 --seed $seed --exp_name_stem $3 --num_eval_runs 10 > gpurun_out/curves/$3_$seed.log 2>&1 &
  done
  wait
  for seed in 1 2 3 4 5; do cp models/$3_$seed/scalars.jsonl gpurun_out/curves/$3_$seed.jsonl; tail -2 gpurun_out/curves/$3_$seed.log | cut -c1-150; done
done
