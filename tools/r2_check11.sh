set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_table.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_table.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_table.log
tail -n 4 gpurun_out/pytest_table.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"; tail -n 3 gpurun_out/bench1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench1.json').read().strip().splitlines()[-1])
print(d['value'], d['phase_ms'], d['roofline']['frac'], d['e2e']['value'], d['gpu_launches'])
PY
KH_TRACE=1 timeout 300 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline 2>&1 | grep "kh-trace" | tail -n 12
exit 0
