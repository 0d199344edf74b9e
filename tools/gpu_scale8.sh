mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
KH_PARITY_N0=4500000 KH_PARITY_SHORT=1 timeout 400 $TR tools/dist_parity.py 2>&1 | grep "^rank" | sort | tail -n 8; echo "parity rc=${PIPESTATUS[0]}"
KH_TRACE=1 timeout 420 $TR bench.py --gpus $N --steps 1 --warmup 2 2>&1 | grep "kh-trace" | grep -v rebuild | tail -n 24
timeout 420 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench$N rc=$?"
python -c "import json; d=json.loads(open('gpurun_out/bench_${N}gpu.json').read().strip().splitlines()[-1]); print(d['value']/1e9, d['phase_ms'], d['e2e'])"
timeout 420 $TR tools/bench_configs.py counter --steps 2 > gpurun_out/cfg4_counter_${N}gpu.jsonl 2> gpurun_out/cfg4_counter_${N}gpu.err; echo "counter$N rc=$?"
tail -n 2 gpurun_out/cfg4_counter_${N}gpu.jsonl
exit 0
