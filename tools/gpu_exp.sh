KH_EXP_NOATOM=1 KH_TRACE=1 timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 1 --warmup 1 2>&1 | grep -E "scatter|tile.insert" | tail -n 4
exit 0
