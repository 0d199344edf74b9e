"""Single-GPU timing of the multi-GPU sender kernel (scatter_slabs_kernel): p ranks emulated on one GPU
(khx_peer_set maps plain device pointers), so ncu can capture it without NCCL.  usage: slab_probe.py [p] [n]"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch

import kmerhash_b200 as kh
from kmerhash_b200._lib import lib
from bench import generate_workload

p = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000_000
keys, _ = generate_workload(n, 10, seed=23)
d = torch.from_numpy(keys.view(np.int64)).cuda()
t = kh.HashMap(1, "MURMUR64avx", "COUNT")
slab = int(n / p * 1.25) + 4096
peers = [lib.khx_peer_create(0, r, p, 2 * p * slab, 1) for r in range(p)]
for r in range(p):
    for q in range(p):
        if q != r:
            assert lib.khx_peer_set(peers[r], q, lib.khx_peer_recv_ptr(peers[q])) == 0
for it in range(4):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    meta = t.plan_scatter_slabs(d, peers[0], p, slab, it & 1)
    e1.record(); torch.cuda.synchronize()
    print("p=%d n=%d plan_scatter_slabs call %.3f ms (includes launch + sync)" % (p, n, e0.elapsed_time(e1)), flush=True)
print("counts", meta[:, 0].tolist())
for h in peers:
    lib.khx_peer_destroy(h)
