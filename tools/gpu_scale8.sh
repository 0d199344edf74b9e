set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 420 $TR bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench$N rc=$?"
tail -n 2 gpurun_out/bench_${N}gpu.json
timeout 420 $TR tools/bench_configs.py counter --steps 2 > gpurun_out/cfg4_counter_${N}gpu.jsonl 2> gpurun_out/cfg4_counter_${N}gpu.err; echo "counter$N rc=$?"
tail -n 2 gpurun_out/cfg4_counter_${N}gpu.jsonl
tail -n 5 gpurun_out/bench_${N}gpu.err gpurun_out/cfg4_counter_${N}gpu.err
exit 0
