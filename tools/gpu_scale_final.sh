mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 420 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench$N rc=$?"
tail -n 1 gpurun_out/bench_${N}gpu.json | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print(d['value']/1e9, d['finds_per_s']/1e9, d['phase_ms'], d['rank0_steps'], d['e2e'])"
if [ "$N" = "8" ]; then
timeout 420 $TR tools/bench_configs.py counter --steps 2 > gpurun_out/cfg4_counter_${N}gpu.jsonl 2> gpurun_out/cfg4_counter_${N}gpu.err; echo "counter$N rc=$?"
tail -n 1 gpurun_out/cfg4_counter_${N}gpu.jsonl
fi
exit 0
