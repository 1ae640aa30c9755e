cd /root/repo
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --num-envs $((65536 / N)) > gpurun_out/bench_strong_g${N}.json 2> gpurun_out/bench_strong_g${N}.err; echo "strong g$N rc=$?"; tail -1 gpurun_out/bench_strong_g${N}.json | cut -c1-200
