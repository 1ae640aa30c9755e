N=8
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N --steps 2 --warmup 3 --no-cpu-baseline 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['value']), d['phase_ms'])"; }
NCCL_ALGO=Tree run tree 29521
NCCL_ALGO=NVLS run nvls 29522
NCCL_PROTO=LL run ll 29523
NCCL_ALGO=Ring NCCL_PROTO=LL128 run ring_ll128 29524
