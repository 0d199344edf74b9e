mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 4
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_q.json 2>gpurun_out/bench_q.err; echo "rc=$?"; tail -n 3 gpurun_out/bench_q.err
tail -n 1 gpurun_out/bench_q.json | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print(d['value']/1e9, d['phase_ms'], d['roofline']['frac'], {k[:20]:round(v['ms_per_step'],3) for k,v in d['roofline']['other_kernels'].items()}, d['e2e']['value']/1e9, d['e2e']['phase_ms'], d['clocks'])"
exit 0
