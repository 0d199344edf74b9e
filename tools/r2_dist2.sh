set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 900 python -m pytest tests/test_gpu_driver.py -m gpu -q -x --timeout 800 -k "torchrun" > gpurun_out/pytest_dist.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dist.log
tail -n 25 gpurun_out/pytest_dist.log
KH_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 2 --no-e2e > gpurun_out/bench2_trace.json 2> gpurun_out/bench2_trace.err; echo "bench2 trace rc=$?"
grep "kh-trace" gpurun_out/bench2_trace.err | tail -n 70
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench2 rc=$?"
tail -n 5 gpurun_out/bench2.err
cat gpurun_out/bench2.json
exit 0
