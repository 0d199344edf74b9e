"""Experiment (2 GPUs, torchrun): does the local windowed insert slow down once NCCL is up / on received buffers?
Run with KH_TRACE=1; prints stage markers on rank 0."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
import torch.distributed as dist

import kmerhash_b200 as kh
from bench import generate_workload

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
keys, _ = generate_workload(100_000_000, 10, seed=23 + rank)
d = torch.from_numpy(keys.view(np.int64)).cuda()
m = kh.HashMap(1, "MURMUR64avx", "COUNT", device=local, min_load=0.35, max_load=0.8)


def mark(s):
    torch.cuda.synchronize()
    if rank == 0:
        sys.stderr.write("[kh-trace] ===== %s\n" % s)


def once(x):
    m.clear(); m.rehash(128); m.insert(x, estimate=True)


for _ in range(2):
    once(d)
mark("before NCCL init (local keys)")
once(d)
os.environ["NCCL_DEBUG"] = "WARN"
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dist.barrier()
mark("after NCCL init, local keys")
once(d)
out = torch.empty_like(d)
half = d.shape[0] // 2
dist.all_to_all_single(out, d, output_split_sizes=[half, half], input_split_sizes=[half, half])
torch.cuda.synchronize()
mark("after a2a, local keys")
once(d)
mark("after a2a, received buffer")
once(out)
mark("clone of received buffer")
once(out.clone())
dist.barrier()
dist.destroy_process_group()
