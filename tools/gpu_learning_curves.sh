# Learning-curve evidence with the reference's CleanRL defaults (1 env, 2048 steps, 32 minibatches x 10 epochs, 1M steps):
# compare with the reference's published TensorBoard screenshots (HalfCheetah ~1400-1500, Humanoid ~600-650 @ 1M).
set -x
mkdir -p gpurun_out/curves
for seed in 1 2 3; do
  ./build/ppo_continuous_action --env_id HalfCheetah-v5 --seed $seed --exp_name_stem curve_hc --num_eval_runs 1 > gpurun_out/curves/hc_$seed.log 2>&1 &
done
wait
for seed in 1 2 3; do
  ./build/ppo_continuous_action --env_id Humanoid-v4 --seed $seed --exp_name_stem curve_hum --num_eval_runs 1 > gpurun_out/curves/hum_$seed.log 2>&1 &
done
wait
for seed in 1 2; do
  ./build/ppo_continuous_action --env_id Hopper-v5 --seed $seed --exp_name_stem curve_hop --num_eval_runs 1 > gpurun_out/curves/hop_$seed.log 2>&1 &
  ./build/ppo_continuous_action --env_id Ant-v5 --seed $seed --exp_name_stem curve_ant --num_eval_runs 1 > gpurun_out/curves/ant_$seed.log 2>&1 &
done
wait
for f in curve_hc_1 curve_hc_2 curve_hc_3 curve_hum_1 curve_hum_2 curve_hum_3 curve_hop_1 curve_hop_2 curve_ant_1 curve_ant_2; do cp models/$f/scalars.jsonl gpurun_out/curves/$f.jsonl; done
tail -3 gpurun_out/curves/hc_1.log gpurun_out/curves/hum_1.log
