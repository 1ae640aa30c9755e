cd /root/repo
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 600 $TR bench.py --gpus $N --workload ac --steps 10 --warmup 3 > gpurun_out/r02_bench_ac_g${N}_fork.json 2> gpurun_out/r02_bench_ac_g${N}_fork.err; echo "ac g$N rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r02_bench_ac_g${N}_fork.json').read().strip().splitlines()[-1]); print(round(d['value']), d['phase_ms'], d['replicas_bit_identical'])"
