mkdir -p gpurun_out
nvidia-smi -L | head -3
timeout 900 python -m pytest tests/test_gpu_ac_multirank.py tests/test_gpu_peer_exchange.py "tests/test_gpu_edge_cases.py::test_trainers_on_two_devices_in_one_process" -m gpu -q -x > gpurun_out/r02_pytest_2gpu.log 2>&1
tail -6 gpurun_out/r02_pytest_2gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 5 > gpurun_out/r02_bench_g2.json 2> gpurun_out/r02_bench_g2.err
tail -2 gpurun_out/r02_bench_g2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --workload ac --gpus 2 --steps 5 --warmup 3 > gpurun_out/r02_bench_ac_g2.json 2> gpurun_out/r02_bench_ac_g2.err
tail -2 gpurun_out/r02_bench_ac_g2.err
python bench.py --workload ac --steps 5 --warmup 3 > gpurun_out/r02_bench_ac_g1.json 2> gpurun_out/r02_bench_ac_g1.err
python - <<'PY'
import json
for f in ("r02_bench_g2","r02_bench_ac_g2","r02_bench_ac_g1"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, {k:d.get(k) for k in ("value","ms_per_step","phase_ms","exchange_check","replicas_bit_identical","gradient_exchange")}, d["e2e"]["value"], (d.get("cpu_baseline") or {}).get("value"), d["roofline"].get("frac"))
    except Exception as e:
        print(f, "ERR", e)
PY
