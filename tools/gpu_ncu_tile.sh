mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'tile_insert_kernel|region_scatter_kernel' -c 2 -o gpurun_out/full_fused python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/ncu_tile.log 2>&1; echo "ncu rc=$?"
exit 0
