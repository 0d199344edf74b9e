set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_table.py tests/test_gpu_driver.py -m gpu -x -q --timeout 600 -k "level2 or dist_hash_tables or fused or windowed or multi_batch" > gpurun_out/pytest_new.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_new.log
tail -n 12 gpurun_out/pytest_new.log
timeout 900 python bench.py --config 5w --steps 3 --warmup 1 > gpurun_out/bench_5w_1gpu.json 2> gpurun_out/bench_5w_1gpu.err; echo "5w rc=$?"; tail -n 5 gpurun_out/bench_5w_1gpu.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_5w_1gpu.json'))
print(d['value'], d['phase_ms'])
for k,v in d['roofline']['kernels'].items(): print(k[:40], v['ms_per_step'])
PY
KH_TRACE=1 timeout 600 python tools/bench_configs.py k63 --steps 1 > gpurun_out/config3_trace.jsonl 2> gpurun_out/config3_trace.err; grep "kh-trace" gpurun_out/config3_trace.err | head -n 40
exit 0
