set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_table.py -m gpu -q -x --timeout 300 -k "slab or capacity or peer or exchange or query" > gpurun_out/pytest_new.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_new.log
tail -n 15 gpurun_out/pytest_new.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"
tail -n 5 gpurun_out/bench1.err
cat gpurun_out/bench1.json
exit 0
