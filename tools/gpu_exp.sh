mkdir -p gpurun_out
python tools/slab_probe.py 2 2>&1 | tail -n 3
python tools/slab_probe.py 8 2>&1 | tail -n 3
timeout 600 ncu --set full --clock-control none --import-source on -k regex:plan_scatter_slabs -c 2 -o gpurun_out/full_slabs python tools/slab_probe.py 8 > gpurun_out/ncu_slabs.log 2>&1; echo "ncu rc=$?"
exit 0
