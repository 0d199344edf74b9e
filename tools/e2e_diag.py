#!/usr/bin/env python
"""per-step wall times of kh_insert_host on pinned buffers (bench.py's e2e leg, one phase at a time)"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kmerhash_b200 as kh
from bench import generate_workload, pinned_u64

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
keys, q = generate_workload(n, n // 10, seed=23)
hk = pinned_u64(keys); hk_np = hk.numpy().view(np.uint64)
d = torch.empty(n, dtype=torch.int64, device="cuda")
for _ in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter(); d.copy_(hk, non_blocking=True); torch.cuda.synchronize()
    print("plain H2D %.2f ms" % (1e3 * (time.perf_counter() - t0)))
m = kh.HashMap(1, "MURMUR64avx", "COUNT", "ROBINHOOD_OFFSETS", device=0, min_load=0.35, max_load=0.8)
for i in range(6):
    m.clear(); m.rehash(128)
    t0 = time.perf_counter(); m.insert(hk_np, estimate=True); t1 = time.perf_counter()
    print("insert_host step %d: %.2f ms" % (i, 1e3 * (t1 - t0)), flush=True)
for i in range(3):
    m.clear(); m.rehash(128)
    torch.cuda.synchronize(); t0 = time.perf_counter(); m.insert(d, estimate=True); torch.cuda.synchronize(); t1 = time.perf_counter()
    print("insert_dev step %d: %.2f ms" % (i, 1e3 * (t1 - t0)), flush=True)
