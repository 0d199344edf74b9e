set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_table.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_table.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_table.log
tail -n 5 gpurun_out/pytest_table.log
KH_TRACE=1 timeout 600 python bench.py --steps 3 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/bench_trace.json 2> gpurun_out/bench_trace.err; echo "bench rc=$?"
grep -E "level1|tiles|total" gpurun_out/bench_trace.err | tail -n 9
timeout 600 python bench.py --steps 5 --warmup 2 --no-cpu-baseline > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err; echo "bench rc=$?"
cat gpurun_out/bench_plain.json
exit 0
