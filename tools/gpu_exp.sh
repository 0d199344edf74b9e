python tools/slab_probe.py 2 2>&1 | tail -n 2
python tools/slab_probe.py 4 2>&1 | tail -n 2
python tools/slab_probe.py 8 2>&1 | tail -n 2
exit 0
