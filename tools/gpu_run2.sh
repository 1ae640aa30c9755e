set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for S in 8 16 32; do PB_SPLITS=$S python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('S$S', d['value'], d['e2e']['value'], d['phase_ms'])"; done
PB_SPLITS=16 python tools/profile_update.py 1 32 > gpurun_out/plain.log 2>&1 && PB_SPLITS=16 ncu --set full --import-source on --clock-control none -k regex:"fwd_bwd|wgrad" -s 40 -c 2 -f -o gpurun_out/r01_update_v2 python tools/profile_update.py 1 32 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
