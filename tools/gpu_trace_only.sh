N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
KH_TRACE=1 timeout 300 $TR bench.py --gpus $N --steps 1 --warmup 2 --no-e2e 2>&1 | grep "kh-trace" | grep "dist\.\|tile.scatter" | tail -n 30
exit 0
