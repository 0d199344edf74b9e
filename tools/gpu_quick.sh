set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo "bench rc=$?"
cat gpurun_out/bench_q.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['phase_ms'], d['roofline']['kernel_ms_per_step'], d['e2e']['value']/1e9)"
KH_TRACE=1 timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 1 --warmup 1 2>&1 | grep -v "^{" | tail -n 40
KH_TILE=0 timeout 600 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/bench_q_notile.json 2>&1
exit 0
