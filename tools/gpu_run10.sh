timeout 600 python -m pytest tests -m gpu -q -k "ppo or golden or cli" 2>&1 | tail -3
PB_L1_DEBUG=2 timeout 300 python tools/profile_update.py 2 32 2>&1 | tail -3
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['phase_ms'], d['gpu_launches'])"
