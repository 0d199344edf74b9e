set -x
mkdir -p gpurun_out
timeout 300 python tools/kmergen_bench.py 2000000 > gpurun_out/plain_kmer.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'kmer_tiles_kernel' -s 1 -c 1 -o gpurun_out/r02c_kmergen python tools/kmergen_bench.py 2000000 > gpurun_out/ncu_kmer.log 2>&1; echo "ncu rc=$?"
tail -n 3 gpurun_out/ncu_kmer.log
exit 0
