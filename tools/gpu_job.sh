cd /root/repo
timeout 900 python -m pytest tests/test_gpu_ac_parity.py tests/test_cli.py -x -q -m gpu 2>&1 | tail -2
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 600 python bench.py --workload ac --steps 10 --warmup 3 > gpurun_out/r02_bench_ac_final.json 2> gpurun_out/r02_bench_ac_final.err; echo "bench ac rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r02_bench_ac_final.json').read().strip().splitlines()[-1]); print(round(d['value']), d['ms_per_step'], d['phase_ms'])"
