mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_table.py -m gpu -q -x -k "slab or window or tile or fused or multi_gpu" 2>&1 | tail -n 3
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
KH_PARITY_SHORT=1 KH_PARITY_N0=5000000 timeout 600 $TR tools/dist_parity.py 2>&1 | grep "^rank" | sort | tail -n 8; echo "parity rc=${PIPESTATUS[0]}"
timeout 420 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench$N rc=$?"
tail -n 1 gpurun_out/bench_${N}gpu.json | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print(d['value']/1e9, d['finds_per_s']/1e9, d['finds_per_s_median']/1e9, d['phase_ms'], d['e2e']['value']/1e9)"
exit 0
