set -x
for TM in 32 64; do for OCC in 2 3; do PB_FB_TM=$TM PB_WG_OCC=$OCC python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('TM$TM OCC$OCC', d['value'], d['e2e']['value'], d['phase_ms'])"; done; done
