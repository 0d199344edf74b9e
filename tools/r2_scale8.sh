set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -n 8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench8.json 2> gpurun_out/bench8.err; echo "bench8 rc=$?"; tail -n 3 gpurun_out/bench8.err
tail -n 1 gpurun_out/bench8.json | cut -c1-1200
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --config 5w --steps 3 --warmup 1 --no-parity > gpurun_out/bench8_5w.json 2> gpurun_out/bench8_5w.err; echo "bench8 5w rc=$?"; tail -n 3 gpurun_out/bench8_5w.err
tail -n 1 gpurun_out/bench8_5w.json | cut -c1-1200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 tools/bench_configs.py counter > gpurun_out/config4_8gpu_fused.jsonl 2> gpurun_out/config4_8.err; tail -n 3 gpurun_out/config4_8.err; cat gpurun_out/config4_8gpu_fused.jsonl
KH_COUNTER_FUSED=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29524 tools/bench_configs.py counter > gpurun_out/config4_8gpu_unfused.jsonl 2>> gpurun_out/config4_8.err; cat gpurun_out/config4_8gpu_unfused.jsonl
exit 0
