cd /root/repo
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513"
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02_bench_g8.json 2> gpurun_out/r02_bench_g8.err; echo "ppo g8 rc=$?"; tail -1 gpurun_out/r02_bench_g8.json | cut -c1-200
timeout 600 $TR bench.py --gpus 8 --workload ac --steps 10 --warmup 3 > gpurun_out/r02_bench_ac_g8.json 2> gpurun_out/r02_bench_ac_g8.err; echo "ac g8 rc=$?"; tail -1 gpurun_out/r02_bench_ac_g8.json | cut -c1-200
PB_PEER_AR=0 timeout 600 $TR bench.py --gpus 8 --steps 10 --warmup 5 > gpurun_out/r02_bench_g8_nccl.json 2> gpurun_out/r02_bench_g8_nccl.err; echo "ppo g8 nccl rc=$?"; tail -1 gpurun_out/r02_bench_g8_nccl.json | cut -c1-200
