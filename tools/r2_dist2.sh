set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hash_kmer.py tests/test_gpu_table.py tests/test_gpu_driver.py -m gpu -q -x --timeout 800 -k "kmers or pipelined or torchrun or barrier" > gpurun_out/pytest_dist.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dist.log
tail -n 25 gpurun_out/pytest_dist.log
timeout 300 python tools/kmergen_bench.py > gpurun_out/kmergen.jsonl 2> gpurun_out/kmergen.err; cat gpurun_out/kmergen.jsonl
KH_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 2 --no-e2e --no-parity > gpurun_out/bench2_trace.json 2> gpurun_out/bench2_trace.err; echo "bench2 trace rc=$?"
grep "kh-trace" gpurun_out/bench2_trace.err | grep -v "rebuild  " | tail -n 40
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench2 rc=$?"
tail -n 5 gpurun_out/bench2.err
cat gpurun_out/bench2.json
KH_PIPE_CHUNKS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e --no-parity > gpurun_out/bench2_nopipe.json 2> gpurun_out/bench2_nopipe.err; echo "bench2 nopipe rc=$?"
cat gpurun_out/bench2_nopipe.json | cut -c1-400
KH_PIPE_CHUNKS=8 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e --no-parity > gpurun_out/bench2_pipe8.json 2> gpurun_out/bench2_pipe8.err; echo "bench2 pipe8 rc=$?"
cat gpurun_out/bench2_pipe8.json | cut -c1-400
exit 0
