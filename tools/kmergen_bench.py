#!/usr/bin/env python
"""stand-alone k-mer generation rate (150/120 B of bases read + 8 B written = 9.25 B per 31-mer)"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import kmerhash_b200 as kh
from kmerhash_b200 import ops
from tools.bench_configs import synth_reads

n_reads = int(sys.argv[1]) if len(sys.argv) > 1 else 8_333_334
peak = 6553.3
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
bases, offs = synth_reads(n_reads, 150, 100_000_000, 0)
for k, canon in ((31, True), (31, False), (63, True), (21, True)):
    ts = []
    for it in range(4):
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); out = ops.kmers_from_reads_dev(bases, offs, k, canon); e1.record(); torch.cuda.synchronize()
        if it:
            ts.append(e0.elapsed_time(e1))
        n = out.shape[0]
        del out
    ms = sum(ts) / len(ts)
    W = 1 if k <= 32 else 2
    byt = bases.shape[0] + 8 * W * n
    print(json.dumps({"k": k, "canonical": canon, "reads": n_reads, "kmers": n, "ms": ms, "kmers_per_s": n / ms * 1e3,
                      "algorithmic_GBps": byt / ms / 1e6, "frac_of_hbm_peak": byt / ms / 1e6 / peak}), flush=True)
