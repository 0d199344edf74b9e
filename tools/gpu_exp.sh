N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for i in 1 2; do
timeout 420 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-e2e 2>/dev/null | tail -n 1 | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print(d['value']/1e9, d['finds_per_s']/1e9, d['phase_ms'], d['rank0_steps'])"
done
exit 0
