set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_table.py -m gpu -x -q --timeout 300 > gpurun_out/pytest_table.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_table.log
tail -n 8 gpurun_out/pytest_table.log
for big in 1 0; do
KH_L1_BIG=$big timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/bench_big$big.json 2> gpurun_out/bench_big.err; python -c "
import json
d=json.load(open('gpurun_out/bench_big$big.json')); print('l1_big $big', round(d['value']/1e9,2), d['phase_ms']['insert'], [round(v['ms_per_step'],3) for v in d['roofline']['kernels'].values()])"
done
KH_L1_BIG=1 timeout 600 python tools/bench_configs.py k63 --steps 2 --n 100000000 2> gpurun_out/c3.err | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(d['hash'], round(d['inserts_per_s']/1e9,2), d['phase_ms'])" | head -n 2
exit 0
