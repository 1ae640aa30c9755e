cd /root/repo
timeout 900 python -m pytest tests/test_gpu_ac_multirank.py tests/test_gpu_ac_parity.py tests/test_gpu_env_parity.py tests/test_gpu_trained_state_parity.py -x -q -m gpu 2>&1 | tail -3
N=$(nvidia-smi -L | wc -l)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
timeout 600 $TR bench.py --gpus $N --workload ac --steps 10 --warmup 3 > gpurun_out/r02_bench_ac_g${N}_ll.json 2> gpurun_out/r02_bench_ac_g${N}_ll.err; echo "ac g$N rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r02_bench_ac_g${N}_ll.json').read().strip().splitlines()[-1]); print(round(d['value']), d['phase_ms'], d['replicas_bit_identical'], d['roofline']['traffic'])"
