# Weak-scaling bench lines of both workloads on all GPUs of the box: gpurun --gpus N -- 'bash tools/gpu_scaling.sh'
# (strong scaling of the ppo workload: add STRONG=1, which fixes the total at 65 536 envs)
cd /root/repo
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_g${N}.json 2> gpurun_out/bench_g${N}.err; echo "ppo g$N rc=$?"; tail -1 gpurun_out/bench_g${N}.json | cut -c1-200
timeout 600 $TR bench.py --gpus $N --workload ac --steps 10 --warmup 3 > gpurun_out/bench_ac_g${N}.json 2> gpurun_out/bench_ac_g${N}.err; echo "ac g$N rc=$?"; tail -1 gpurun_out/bench_ac_g${N}.json | cut -c1-200
if [ -n "$STRONG" ]; then
  timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --num-envs $((65536 / N)) > gpurun_out/bench_strong_g${N}.json 2> gpurun_out/bench_strong_g${N}.err; echo "strong g$N rc=$?"; tail -1 gpurun_out/bench_strong_g${N}.json | cut -c1-200
fi
