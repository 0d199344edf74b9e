mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m pytest tests/test_gpu_table.py -m gpu -x -q -k "slab or fused or multi_gpu or barrier" 2>&1 | tail -n 5
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tools/dist_parity.py 2>&1 | grep -v "^W\|^\*\*\*" | tail -n 12; echo "parity rc=${PIPESTATUS[0]}"
KH_TRACE=1 timeout 420 $TR bench.py --gpus $N --steps 1 --warmup 2 --no-e2e 2>&1 | grep "kh-trace" | grep "dist\." | tail -n 14
timeout 420 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench$N rc=$?"
python -c "import json; d=json.loads(open('gpurun_out/bench_${N}gpu.json').read().strip().splitlines()[-1]); print(d['value']/1e9, d['finds_per_s']/1e9, d['phase_ms'], d['e2e'])"
exit 0
