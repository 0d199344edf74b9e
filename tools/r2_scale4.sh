set -x
mkdir -p gpurun_out
N=${1:-4}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for cfg in "1 sm" "4 ce"; do
  set -- $cfg
  KH_PIPE_CHUNKS=$1 KH_PIPE_MODE=$2 timeout 600 $TR --master-port 2953$1 bench.py --gpus $N --steps 8 --warmup 3 --no-e2e --no-parity > gpurun_out/bench${N}_$1_$2.json 2> gpurun_out/bench${N}_$1_$2.err; echo "bench$N $cfg rc=$?"
  tail -n 1 gpurun_out/bench${N}_$1_$2.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pieces', d['config'].get('exchange_pieces'), round(d['value']/1e9,1), d['phase_ms'])"
done
KH_TRACE=1 KH_PIPE_CHUNKS=4 KH_PIPE_MODE=ce timeout 300 $TR --master-port 29539 bench.py --gpus $N --steps 2 --warmup 2 --no-e2e --no-parity 2>&1 | grep "kh-trace\] pipelined(ce, rank [01])\|bulk\.\|insert.total" | tail -n 14
exit 0
