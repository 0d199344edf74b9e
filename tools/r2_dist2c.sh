set -x
mkdir -p gpurun_out
timeout 120 tools/peer_write_bw > gpurun_out/peer_write_bw.jsonl 2>&1; tail -n 2 gpurun_out/peer_write_bw.jsonl
timeout 900 python -m pytest tests/test_gpu_table.py tests/test_gpu_driver.py -m gpu -q -x --timeout 800 -k "fused_with_sender or torchrun or slab or exchange or query" > gpurun_out/pytest_dist.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dist.log
tail -n 4 gpurun_out/pytest_dist.log
python tools/sender_local_bench.py > gpurun_out/sender_local.json; cat gpurun_out/sender_local.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench2 rc=$?"
tail -n 3 gpurun_out/bench2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench2.json').read().strip().splitlines()[-1])
print(d['value'], d['phase_ms'], d.get('e2e'))
PY
bash tools/gpu_trace_only.sh 2 > gpurun_out/trace2.txt 2>&1; tail -n 14 gpurun_out/trace2.txt
exit 0
