set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29531 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench8_default.json 2> gpurun_out/bench8_default.err; echo "bench8 default rc=$?"
tail -n 1 gpurun_out/bench8_default.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pieces', d['config'].get('exchange_pieces'), round(d['value']/1e9,1), d['phase_ms'], d.get('e2e'))"
KH_PIPE_CHUNKS=8 timeout 600 $TR --master-port 29532 bench.py --gpus 8 --steps 8 --warmup 3 --no-e2e --no-parity > gpurun_out/bench8_8_ce.json 2> gpurun_out/bench8_8_ce.err; echo "bench8 8 ce rc=$?"
tail -n 1 gpurun_out/bench8_8_ce.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pieces', d['config'].get('exchange_pieces'), round(d['value']/1e9,1), d['phase_ms'])"
KH_TRACE=1 timeout 300 $TR --master-port 29539 bench.py --gpus 8 --steps 2 --warmup 2 --no-e2e --no-parity 2>&1 | grep "kh-trace\] pipelined(ce, rank [07])" | tail -n 4
exit 0
