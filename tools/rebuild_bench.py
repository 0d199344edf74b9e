#!/usr/bin/env python
"""tile_rebuild_kernel alone: a table of --distinct keys rehashed up and down (grow x2, shrink x2), kernel time from the
table's own events.   usage: python tools/rebuild_bench.py [--distinct 18000000] [--w 1]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import kmerhash_b200 as kh

ap = argparse.ArgumentParser()
ap.add_argument("--distinct", type=int, default=18_000_000)
ap.add_argument("--w", type=int, default=1)
ap.add_argument("--hash", default="MURMUR64avx")
a = ap.parse_args()
W = a.w
g = torch.Generator(device="cuda"); g.manual_seed(5)
keys = torch.randint(0, 2**62, (a.distinct, W) if W == 2 else (a.distinct,), generator=g, device="cuda", dtype=torch.int64)
m = kh.HashMap(W, a.hash, "COUNT", min_load=0.05, max_load=0.8)
m.insert(keys)
size, cap0 = m.size(), m.capacity()
m.kernel_timing(True)
out = []
for it in range(3):
    for cap in (cap0 * 2, cap0):
        ms0, n0 = m.kernel_time("rebuild")
        m.rehash(cap)
        ms1, n1 = m.kernel_time("rebuild")
        assert m.capacity() == cap and m.size() == size
        out.append({"from": cap0 if cap != cap0 else cap0 * 2, "to": cap, "ms": round(ms1 - ms0, 4), "launches": n1 - n0})
slot = 16 * W
for o in out:
    o["GBps"] = round((o["from"] + o["to"]) * slot / o["ms"] / 1e6, 1)
print(json.dumps({"size": size, "W": W, "hash": a.hash, "rebuilds": out}))
assert int(m.count(keys[:1000000]).sum()) == 1000000
