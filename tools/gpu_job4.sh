cd /root/repo
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512"
timeout 600 $TR bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r02_bench_g4.json 2> gpurun_out/r02_bench_g4.err; echo "ppo g4 rc=$?"; tail -1 gpurun_out/r02_bench_g4.json | cut -c1-200
timeout 600 $TR bench.py --gpus 4 --steps 5 --warmup 3 --num-envs 16384 > gpurun_out/r02_bench_strong_g4.json 2> gpurun_out/r02_bench_strong_g4.err; echo "strong g4 rc=$?"; tail -1 gpurun_out/r02_bench_strong_g4.json | cut -c1-200
timeout 600 $TR bench.py --gpus 4 --workload ac --steps 10 --warmup 3 > gpurun_out/r02_bench_ac_g4.json 2> gpurun_out/r02_bench_ac_g4.err; echo "ac g4 rc=$?"; tail -1 gpurun_out/r02_bench_ac_g4.json | cut -c1-200
