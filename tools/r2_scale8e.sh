set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 100 $TR --master-port 29531 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench8_default.json 2> gpurun_out/bench8_default.err; echo "bench8 default rc=$?"
tail -n 1 gpurun_out/bench8_default.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pieces', d['config'].get('exchange_pieces'), round(d['value']/1e9,1), d['phase_ms'], d.get('e2e'))"
KH_PIPE_CHUNKS=1 timeout 60 $TR --master-port 29532 bench.py --gpus 8 --steps 8 --warmup 3 --no-e2e --no-parity > gpurun_out/bench8_1_sm.json 2> gpurun_out/bench8_1_sm.err; echo "bench8 1 sm rc=$?"
tail -n 1 gpurun_out/bench8_1_sm.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pieces', d['config'].get('exchange_pieces'), round(d['value']/1e9,1), d['phase_ms'])"
exit 0
