mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
KH_TRACE=1 timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 1 --warmup 1 2>&1 | grep -v "^{" | tail -n 14
for m in old new; do
KH_SCATTER=$m timeout 600 python bench.py --no-cpu-baseline --no-e2e > gpurun_out/bench_q_$m.json 2> gpurun_out/bench_q.err; echo "bench $m rc=$?"
cat gpurun_out/bench_q_$m.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['phase_ms'], d['roofline']['kernel_ms_per_step'], {k[:22]:v['ms_per_step'] for k,v in d['roofline']['other_kernels'].items()})"
done
exit 0
