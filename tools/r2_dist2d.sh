set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_driver.py -m gpu -q -x --timeout 800 -k "torchrun" > gpurun_out/pytest_dist.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dist.log
tail -n 6 gpurun_out/pytest_dist.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for cfg in "1 sm" "4 sm" "4 ce" "2 ce" "8 ce"; do
  set -- $cfg
  KH_PIPE_CHUNKS=$1 KH_PIPE_MODE=$2 timeout 600 $TR --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e > gpurun_out/bench2_$1_$2.json 2> gpurun_out/bench2.err; echo "bench2 $cfg rc=$?"
  tail -n 2 gpurun_out/bench2.err | cut -c1-300
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench2_$1_$2.json').read().strip().splitlines()[-1])
print("$cfg", d['value'], d['phase_ms'])
PY
done
exit 0
