set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.used,memory.total --format=csv
timeout 1200 python -m pytest tests/test_gpu_table.py -m gpu -q -x --timeout 300 > gpurun_out/pytest_table.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_table.log
tail -n 30 gpurun_out/pytest_table.log
KH_TRACE=1 timeout 600 python bench.py --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_trace.json 2> gpurun_out/bench_trace.err; echo "bench rc=$?"
tail -n 60 gpurun_out/bench_trace.err
cat gpurun_out/bench_trace.json
exit 0
