cd /root/repo
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_g${N}_ll.json 2> gpurun_out/r02_bench_g${N}_ll.err; echo "ppo g$N rc=$?"; tail -1 gpurun_out/r02_bench_g${N}_ll.json | cut -c1-200
timeout 600 $TR bench.py --gpus $N --workload ac --steps 10 --warmup 3 > gpurun_out/r02_bench_ac_g${N}_v3.json 2> gpurun_out/r02_bench_ac_g${N}_v3.err; echo "ac g$N rc=$?"; tail -1 gpurun_out/r02_bench_ac_g${N}_v3.json | cut -c1-200
