set -x
mkdir -p gpurun_out
for pc in 4 2; do
KH_PIPE_CHUNKS=$pc timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2953$pc bench.py --gpus 8 --steps 10 --warmup 3 --no-e2e --no-parity > gpurun_out/bench8_pipe$pc.json 2> gpurun_out/bench8_pipe$pc.err; echo "pipe $pc rc=$?"; tail -n 2 gpurun_out/bench8_pipe$pc.err
tail -n 1 gpurun_out/bench8_pipe$pc.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pipe', d['config']['exchange_pieces'], round(d['value']/1e9,1), d['phase_ms'])"
done
KH_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29539 bench.py --gpus 8 --steps 3 --warmup 2 --no-e2e --no-parity > gpurun_out/bench8_trace.json 2> gpurun_out/bench8_trace.err; grep "kh-trace\] dist" gpurun_out/bench8_trace.err | tail -n 16
exit 0
