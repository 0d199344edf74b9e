"""Measurement helper (GPU box): random-access roofline probes and a bucket-geometry sweep of the
insert / find kernels on the BenchmarkHashTables-shaped workload.  Prints JSON lines."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import numpy as np
import torch

import kmerhash_b200 as kh


def gen_keys(n, repeats=10, seed=23, words=1):
    rng = np.random.default_rng(seed)
    n_uniq = int(n / (1 + (repeats - 1) / 2.0)) + 1
    if words == 1:
        uniq = rng.integers(0, 2**62, size=n_uniq, dtype=np.uint64) & np.uint64(0x7FFFFFFF7FFFFFFF)
    else:
        uniq = rng.integers(0, 2**62, size=(n_uniq, 2), dtype=np.uint64)
    reps = rng.integers(1, repeats + 1, size=n_uniq)
    keys = np.repeat(uniq, reps, axis=0)[:n]
    keys = keys[rng.permutation(len(keys))]
    return np.ascontiguousarray(keys)


def probes():
    for buf in (512 << 20, 4 << 30):
        for mode in ("read", "rmw"):
            for b in (16, 32, 64, 128):
                n = 200_000_000
                ms = kh.probe_random_access(buf, n, b, mode, iters=3)
                print(json.dumps({"probe": "random", "buf_mb": buf >> 20, "mode": mode, "bytes": b,
                                  "ms": round(ms, 3), "G_access_per_s": round(n / ms / 1e6, 2),
                                  "GB_per_s": round(n * b / ms / 1e6, 1)}), flush=True)
    ms = kh.probe_stream_copy(4 << 30, 5)
    print(json.dumps({"probe": "stream_copy", "GB_per_s": round(2 * (4 << 30) / ms / 1e6, 1)}), flush=True)


def l2probes():
    """does a table region that fits the 126 MB L2 raise the random-access ceiling? (partitioned-insert question)"""
    for buf_mb in (16, 32, 64, 96, 128, 256, 512):
        for mode in ("read", "rmw", "read_ilp4", "rmw_ilp4"):
            for b in (32, 64):
                n = 100_000_000
                ms = kh.probe_random_access(buf_mb << 20, n, b, mode, iters=3)
                print(json.dumps({"probe": "random", "buf_mb": buf_mb, "mode": mode, "bytes": b, "ms": round(ms, 3),
                                  "G_access_per_s": round(n / ms / 1e6, 2), "GB_per_s": round(n * b / ms / 1e6, 1)}), flush=True)


def geom_sweep(n=100_000_000, nq=10_000_000, words=1):
    keys = gen_keys(n, words=words)
    d = torch.from_numpy(keys.view(np.int64)).cuda()
    q = d[:nq].contiguous()
    geoms = [(64, 1), (32, 1), (32, 2), (32, 4)] + ([(16, 2), (16, 4), (16, 8)] if words == 1 else [])
    for lb, g in geoms:
        m = kh.HashMap(words, "MURMUR64avx", "COUNT", max_load=0.8, min_load=0.35)
        m.set_geometry(lb, g)
        res = {"geom": "%dx%d" % (lb, g), "words": words, "n": n}
        for it in range(3):
            m.clear()
            m.kernel_timing(True)
            t0 = time.perf_counter(); new = m.insert(d, estimate=True); t1 = time.perf_counter()
            ims, il = m.kernel_time("insert")
            m.kernel_timing(True)
            m.find(q); t2 = time.perf_counter()
            fms, fl = m.kernel_time("lookup")
            m.kernel_timing(True)
            m.count(q)
            cms, _ = m.kernel_time("lookup")
            m.kernel_timing(True)
            m.erase(q)
            ems, _ = m.kernel_time("lookup")
        res.update({"distinct": new, "capacity": m.capacity(), "insert_call_ms": round((t1 - t0) * 1e3, 2),
                    "insert_kernel_ms": round(ims, 3), "insert_launches": il,
                    "insert_Mkeys_s": round(n / ims / 1e3, 1), "find_kernel_ms": round(fms, 3),
                    "find_Mkeys_s": round(nq / fms / 1e3, 1), "count_Mkeys_s": round(nq / cms / 1e3, 1),
                    "erase_Mkeys_s": round(nq / ems / 1e3, 1)})
        print(json.dumps(res), flush=True)
        del m


if __name__ == "__main__":
    what = sys.argv[1:] or ["probes", "geom"]
    if "probes" in what:
        probes()
    if "l2probes" in what:
        l2probes()
    if "geom" in what:
        geom_sweep()
    if "geom2" in what:
        geom_sweep(words=2)
