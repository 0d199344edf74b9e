mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'tile_insert_kernel|region_scatter_kernel|hll_update_kernel' -c 3 -o gpurun_out/full_tile2 python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > gpurun_out/ncu_tile.log 2>&1; echo "ncu rc=$?"
exit 0
