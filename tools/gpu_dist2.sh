mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 5
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"
python -c "import json; d=json.loads(open('gpurun_out/bench_q.json').read().strip().splitlines()[-1]); print(d['value']/1e9, d['phase_ms'], d['e2e']['value']/1e9, d['e2e']['phase_ms'], d['roofline']['frac'])"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tools/dist_parity.py 2>&1 | grep -v "^W\|^\*\*\*" | tail -n 8; echo "parity rc=${PIPESTATUS[0]}"
bash tools/gpu_scale_trace.sh $N
exit 0
