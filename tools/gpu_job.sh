cd /root/repo
timeout 900 python -m pytest tests/test_gpu_peer_exchange.py tests/test_gpu_ac_multirank.py tests/test_gpu_ac_parity.py -x -q -m gpu > gpurun_out/r02_pytest_2gpu_v3.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02_pytest_2gpu_v3.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 10 --warmup 5 --no-cpu-baseline 2>gpurun_out/g2_ll.err | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('LL', round(d['value']), d['phase_ms'], d['exchange_check'])"
PB_PEER_LL=0 timeout 600 $TR bench.py --gpus 2 --steps 10 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('two-kernel', round(d['value']), d['phase_ms'])"
timeout 300 python bench.py --workload ac --steps 5 --warmup 2 --no-cpu-baseline 2>gpurun_out/ac1.err | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ac 1 gpu', round(d['value']), d['phase_ms'], d['roofline']['kernel_ms'], d['roofline_update']['frac'], d['gpu_launches'])"
tail -2 gpurun_out/g2_ll.err gpurun_out/ac1.err
