set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_table.py tests/test_gpu_driver.py -m gpu -x -q --timeout 600 > gpurun_out/pytest_new.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_new.log
tail -n 12 gpurun_out/pytest_new.log
KH_TRACE=1 timeout 600 python tools/bench_configs.py k63 --steps 1 > gpurun_out/config3_trace.jsonl 2> gpurun_out/config3_trace.err; grep "kh-trace" gpurun_out/config3_trace.err | grep -v "rebuild  " | sed -n 8,22p
timeout 600 python tools/bench_configs.py k63 --steps 3 > gpurun_out/config3_k63.jsonl 2> gpurun_out/config3.err; cut -c1-200 gpurun_out/config3_k63.jsonl; python -c "
import json
for l in open('gpurun_out/config3_k63.jsonl'):
    d=json.loads(l); print(d['hash'], round(d['inserts_per_s']/1e9,2), round(d['finds_per_s']/1e9,2), d['phase_ms'])"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench1b.json 2> gpurun_out/bench1b.err; python -c "
import json
d=json.load(open('gpurun_out/bench1b.json')); print(d['value'], d['phase_ms'])"
exit 0
