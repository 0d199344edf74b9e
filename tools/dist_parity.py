"""Parity of DistributedMap on real GPUs (torchrun, NCCL): every rank's table must hold exactly what the reference's rank
of the same index holds -- { (k, c) in the global count map : DistHash_9876543(k) & (p-1) == rank } -- after a sequence
of batches that exercises the single-pass slab exchange, its overflow fallback (skewed owners), the growth of the
slabs (a larger batch) and the return to the slab path.  The oracle (oracle/, CPU) is the checker.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dist_parity.py
Prints one line per rank; exit code 1 on any mismatch."""
import os
import sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

import kmerhash_b200 as kh
from kmerhash_b200.distributed import DistributedMap
import oracle

rank, local, p = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
os.environ.setdefault("NCCL_DEBUG", "WARN")
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
o = oracle.port


def batch_of(seed, n, skew=False):
    """every rank can regenerate every rank's batch (the checker needs the global multiset)"""
    rng = np.random.default_rng(seed)
    if n == 0:
        return np.empty(0, dtype=np.uint64)
    base = rng.integers(0, 2**62, size=max(1000, n // 5), dtype=np.uint64)
    b = base[rng.integers(0, base.shape[0], size=n)]
    if skew:                                      # 70 % of the batch is ONE key: its owner's slab overflows
        b[rng.random(n) < 0.7] = base[0]
    return np.ascontiguousarray(b)


def sort_pairs(k, v):
    order = np.argsort(k, kind="stable")
    return k[order], v[order]


ok = True
SIZES = [int(x) for x in os.environ.get("KH_PARITY_N0", "5000000,200000").split(",")]   # tile-insert sized and small batches
SHORT = os.environ.get("KH_PARITY_SHORT", "0") not in ("", "0")
for n0 in SIZES:
    m = DistributedMap(1, "MURMUR64avx", "MURMUR64avx", "COUNT", device=local)
    ref = o.Map(1, "COUNT")
    plan = [(n0, False), (n0, False), (n0, True), (n0, False), (3 * n0, False), (3 * n0, False), (n0, True), (n0, True), (n0, False)]
    if SHORT:
        plan = [(n0, False), (n0, True), (n0, False), (3 * n0, False), (n0, False)]
    path = []
    for step, (n, skew) in enumerate(plan):
        mine = None
        for g in range(p):
            b = batch_of(1000 * step + g + n0 % 97, n + 1234 * g, skew)
            ref.insert(b, default=1)
            if g == rank:
                mine = b
        m.insert(torch.from_numpy(mine.view(np.int64)).cuda())
        path.append("slab" if (m.slabs and m.slab_misses == 0) else "exact")
        gk, gv = ref.to_vector()
        own = o.rank("MURMUR64avx", gk, p) == rank
        ek, ev = sort_pairs(gk[own], gv[own])
        tk, tv = sort_pairs(*m.to_vector())
        good = np.array_equal(tk, ek) and np.array_equal(tv.astype(np.uint32), ev.astype(np.uint32))
        ok = ok and good
        if not good:
            print("rank %d n0 %d step %d MISMATCH: %d vs %d entries" % (rank, n0, step, len(tk), len(ek)), flush=True)
    # collective calls with nothing to contribute (the reference: every rank calls, also with 0 elements,
    # KH/distributed_batched_robinhood_map.hpp:843,911-917): rank 0 inserts nothing, the last rank asks nothing
    for g in range(p):
        b = batch_of(555 + g, 300000 if g else 0)
        if b.shape[0]:
            ref.insert(b, default=1)
        if g == rank:
            m.insert(torch.from_numpy(b.view(np.int64)).cuda())
    q0 = batch_of(556 + rank, 0 if rank == p - 1 else 70000)
    r0 = m.find(torch.from_numpy(q0.view(np.int64)).cuda(), 0).cpu().numpy().astype(np.uint32)
    ok = ok and r0.shape[0] == q0.shape[0] and (q0.shape[0] == 0 or np.array_equal(r0, ref.find(q0, 0).astype(np.uint32)))
    c0 = m.count(torch.from_numpy(q0.view(np.int64)).cuda()).cpu().numpy()
    ok = ok and c0.shape[0] == q0.shape[0] and (q0.shape[0] == 0 or np.array_equal(c0, ref.count(q0).astype(np.uint8)))
    if not ok:
        print("rank %d n0 %d empty-batch step FAILED" % (rank, n0), flush=True)
    # queries: per-rank batches (half of the keys absent), a skewed batch (slab overflow -> NCCL path), a larger batch
    # (slabs grow), erase, then counts again.  Answers must come back in input order.
    qpath = []
    for qstep, (nq, skew) in enumerate([(100000, False), (100000, True), (400000, False), (400000, False)]):
        q = batch_of(7000 + 10 * qstep + rank, nq + 321 * rank, skew)
        some = batch_of(1000 + rank + n0 % 97, n0 + 1234 * rank)[: nq // 2]      # keys of the second insert batch
        q[: some.shape[0]] = some
        dq = torch.from_numpy(q.view(np.int64)).cuda()
        found = m.find(dq, 0).cpu().numpy().astype(np.uint32)
        qpath.append(m.last_query_path)
        cnt = m.count(dq).cpu().numpy()
        ok = ok and np.array_equal(found, ref.find(q, 0).astype(np.uint32)) and np.array_equal(cnt, ref.count(q).astype(np.uint8))
    for g in range(p):                           # every rank erases its own batch; the checker erases all of them
        e = batch_of(9000 + g, 50000 + 17 * g)
        e[:20000] = batch_of(1000 + g + n0 % 97, n0 + 1234 * g)[:20000]
        ref.erase(e)
        if g == rank:
            m.erase(torch.from_numpy(e.view(np.int64)).cuda())
    gk, gv = ref.to_vector()
    own = o.rank("MURMUR64avx", gk, p) == rank
    ek, ev = sort_pairs(gk[own], gv[own])
    tk, tv = sort_pairs(*m.to_vector())
    ok = ok and np.array_equal(tk, ek) and np.array_equal(tv.astype(np.uint32), ev.astype(np.uint32))
    q = batch_of(9000 + rank, 50000 + 17 * rank)
    ok = ok and np.array_equal(m.count(torch.from_numpy(q.view(np.int64)).cuda()).cpu().numpy(), ref.count(q).astype(np.uint8))
    print("rank %d n0 %d paths %s queries %s size %d %s" % (rank, n0, ",".join(path), ",".join(qpath), m.local_size(),
                                                           "OK" if ok else "FAIL"), flush=True)
    m.close()                                   # collective: unmap imports -> barrier -> free

# reads -> canonical 31-mers generated inside the sender kernel (DistributedMap.insert_reads): per-rank table contents against
# the oracle's k-mers of ALL ranks' reads restricted to this rank's owner rule
if os.environ.get("KH_PARITY_READS", "1") not in ("", "0"):
    def reads_of(g):
        rng = np.random.default_rng(31000 + g)
        genome = np.frombuffer(b"ACGT", dtype=np.uint8)[rng.integers(0, 4, size=200000)]
        L, nr = 150, 40000 + 1000 * g
        st = rng.integers(0, len(genome) - L, size=nr)
        b = genome[st[:, None] + np.arange(L)[None, :]].copy()
        b[rng.random(b.shape) < 0.0005] = ord("N")
        return b.reshape(-1), (np.arange(nr + 1, dtype=np.uint64) * np.uint64(L))
    m = DistributedMap(1, "MURMUR64avx", "MURMUR64avx", "COUNT", device=local)
    ref = o.Map(1, "COUNT")
    for rep in range(2):
        for g in range(p):
            b, off = reads_of(g)
            ref.insert(o.kmers_from_reads(b, off, 31, True), default=1)
        b, off = reads_of(rank)
        m.insert_reads(torch.from_numpy(b).cuda(), torch.from_numpy(off.astype(np.int64)).cuda(), 31, canonical=True)
    gk, gv = ref.to_vector()
    own = o.rank("MURMUR64avx", gk, p) == rank
    ek, ev = sort_pairs(gk[own], gv[own])
    tk, tv = sort_pairs(*m.to_vector())
    good = np.array_equal(tk, ek) and np.array_equal(tv.astype(np.uint32), ev.astype(np.uint32))
    ok = ok and good
    print("rank %d reads->k-mers fused with the sender: %d distinct %s" % (rank, len(tk), "OK" if good else "FAIL"), flush=True)
    m.close()

flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) else 1)
