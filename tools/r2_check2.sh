set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 5 gpurun_out/pytest_gpu.log
timeout 300 python tools/e2e_diag.py > gpurun_out/e2e_diag.log 2>&1; cat gpurun_out/e2e_diag.log
KH_TRACE=1 timeout 300 python tools/e2e_diag.py 30000000 > gpurun_out/e2e_diag_trace.log 2>&1; tail -n 40 gpurun_out/e2e_diag_trace.log
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'l1_scatter_kernel|tile_insert_kernel|tile_rebuild_kernel' -s 3 -c 3 -o gpurun_out/r02a_bulk python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/ncu_bulk.log 2>&1; echo "ncu rc=$?"
tail -n 5 gpurun_out/ncu_bulk.log
exit 0
