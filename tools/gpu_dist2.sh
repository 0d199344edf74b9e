mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_table.py -x -q -m gpu -k "single_pass or fused_scatter or windowed or tile" 2>&1 | tail -n 15
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tools/dist_parity.py 2>&1 | grep -v "^W\|^\*\*\*" | tail -n 30; echo "parity rc=${PIPESTATUS[0]}"
bash tools/gpu_scale_trace.sh $N
exit 0
