set -x
mkdir -p gpurun_out
timeout 900 python tools/bench_configs.py lp > gpurun_out/cfg2_lp.jsonl 2> gpurun_out/cfg2_lp.err; echo "lp rc=$?"
timeout 900 python tools/bench_configs.py k63 > gpurun_out/cfg3_k63.jsonl 2> gpurun_out/cfg3_k63.err; echo "k63 rc=$?"
timeout 900 python tools/bench_configs.py counter > gpurun_out/cfg4_counter_1gpu.jsonl 2> gpurun_out/cfg4_counter.err; echo "counter rc=$?"
tail -3 gpurun_out/cfg*.err
