mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -n 3
timeout 300 python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench1 rc=$?"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
KH_PARITY_SHORT=1 timeout 600 $TR tools/dist_parity.py 2>&1 | grep "^rank" | sort | tail -n 8; echo "parity rc=${PIPESTATUS[0]}"
timeout 420 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench$N rc=$?"
tail -n 1 gpurun_out/bench_${N}gpu.json | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print(d['value']/1e9, d['finds_per_s']/1e9, d['phase_ms'], d['rank0_steps'][0], d['e2e']['value']/1e9)"
exit 0
