set -x
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'l1_scatter_kernel' -s 1 -c 1 -o gpurun_out/r02b_l1 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_l1.log 2>&1; echo "ncu rc=$?"
tail -n 3 gpurun_out/ncu_l1.log
exit 0
