N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
KH_PARITY_SHORT=1 timeout 600 $TR tools/dist_parity.py 2>&1 | grep -v "^W\|^\*\*\*\|OMP_NUM" | tail -n 25; echo "parity rc=${PIPESTATUS[0]}"
exit 0
