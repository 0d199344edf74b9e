#!/usr/bin/env python
"""How much of the multi-GPU sender's time is SM work and how much NVLink: scatter_slabs_kernel for one rank of p, all p receive
buffers on THIS GPU (no link in the way).  Compare with the traced `dist.scatter->slabs` stage of a real p-GPU run.
usage: python tools/sender_local_bench.py [--n 100000000] [--p 8]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
import kmerhash_b200 as kh
from kmerhash_b200._lib import lib

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=100_000_000)
ap.add_argument("--p", type=int, default=8)
ap.add_argument("--w", type=int, default=1)
a = ap.parse_args()
p, n, W = a.p, a.n, a.w
g = torch.Generator(device="cuda"); g.manual_seed(3)
keys = torch.randint(0, 2**62, (n, W) if W == 2 else (n,), generator=g, device="cuda", dtype=torch.int64)
slab_cap = int(n / p * 1.1) + 4096
cap = 2 * p * slab_cap + 64
peers = [lib.khx_peer_create(0, r, p, cap, W) for r in range(p)]
assert all(peers)
for q in range(1, p):
    assert lib.khx_peer_set(peers[0], q, lib.khx_peer_recv_ptr(peers[q])) == 0
t = kh.HashMap(W, "MURMUR64avx", "COUNT")
cnt = torch.zeros(p, dtype=torch.int64, device="cuda")


def sync():
    assert lib.kh_count_dev(t._h, None, 0, None) == 0


times = []
for it in range(8):
    sync(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    t.scatter_slabs(keys, peers[0], slab_cap, it & 1, cnt, reset=True)
    sync()
    times.append((time.perf_counter() - t0) * 1e3)
c = cnt.cpu().numpy()
assert int(c.sum()) == n and int(c.max()) <= slab_cap
med = float(np.median(times[2:]))
print(json.dumps({"sender_local_ms": [round(x, 3) for x in times], "median_ms": med, "n": n, "p": p, "W": W,
                  "bytes_moved": 2 * 8 * W * n, "GBps": 2 * 8 * W * n / med / 1e6}))
for h in peers:
    lib.khx_peer_destroy(h)
