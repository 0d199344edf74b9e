set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 8 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --config 5w --steps 3 --warmup 1 > gpurun_out/bench_5w_1gpu.json 2> gpurun_out/bench_5w_1gpu.err; echo "5w rc=$?"; tail -n 5 gpurun_out/bench_5w_1gpu.err
cat gpurun_out/bench_5w_1gpu.json | cut -c1-3000
timeout 900 python tools/bench_configs.py k63 --steps 2 > gpurun_out/config3_k63.jsonl 2> gpurun_out/config3.err; echo "k63 rc=$?"; tail -n 3 gpurun_out/config3.err; cut -c1-700 gpurun_out/config3_k63.jsonl
timeout 1500 python tools/bench_configs.py lp --steps 2 > gpurun_out/config2_lp.jsonl 2> gpurun_out/config2.err; echo "lp rc=$?"; tail -n 3 gpurun_out/config2.err; cut -c1-500 gpurun_out/config2_lp.jsonl
exit 0
