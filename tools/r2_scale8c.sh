set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
for cfg in "1 sm" "4 ce"; do
  set -- $cfg
  KH_PIPE_CHUNKS=$1 KH_PIPE_MODE=$2 timeout 600 $TR --master-port 2953$1 bench.py --gpus 8 --steps 10 --warmup 3 --no-e2e > gpurun_out/bench8_$1_$2.json 2> gpurun_out/bench8_$1_$2.err; echo "bench8 $cfg rc=$?"
  tail -n 2 gpurun_out/bench8_$1_$2.err | cut -c1-300
  tail -n 1 gpurun_out/bench8_$1_$2.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pieces', d['config'].get('exchange_pieces'), round(d['value']/1e9,1), d['phase_ms'])"
done
KH_TRACE=1 timeout 300 $TR --master-port 29539 bench.py --gpus 8 --steps 2 --warmup 2 --no-e2e --no-parity 2>&1 | grep "kh-trace\] dist" | tail -n 9
exit 0
