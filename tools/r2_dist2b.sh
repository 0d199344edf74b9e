set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_table.py tests/test_gpu_driver.py tests/test_gpu_hash_kmer.py -m gpu -q -x --timeout 800 -k "fused_with_sender or torchrun or dist_hash_tables or golden or hash" > gpurun_out/pytest_dist.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dist.log
tail -n 25 gpurun_out/pytest_dist.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench2 rc=$?"
tail -n 5 gpurun_out/bench2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench2.json'))
print(d['value'], d['phase_ms'], d.get('e2e'))
for st in d['rank0_steps'][0]['find_stages_host_ms']: print(st)
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29515 tools/bench_configs.py counter --reads 2083334 > gpurun_out/config4_2gpu_fused.jsonl 2> gpurun_out/config4.err; tail -n 3 gpurun_out/config4.err; cat gpurun_out/config4_2gpu_fused.jsonl
KH_COUNTER_FUSED=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29516 tools/bench_configs.py counter --reads 2083334 > gpurun_out/config4_2gpu_unfused.jsonl 2>> gpurun_out/config4.err; cat gpurun_out/config4_2gpu_unfused.jsonl
exit 0
