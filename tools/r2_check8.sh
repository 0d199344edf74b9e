set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hash_kmer.py -m gpu -q -x --timeout 600 > gpurun_out/pytest_kmer.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_kmer.log
tail -n 15 gpurun_out/pytest_kmer.log
timeout 300 python tools/kmergen_bench.py 4000000 > gpurun_out/kmergen.jsonl 2> gpurun_out/kmergen.err; cat gpurun_out/kmergen.jsonl
exit 0
