"""World-size-2 (and 3) gloo tests of the one-process-per-GPU exchange logic (kmerhash_b200/distributed.py).

The CUDA engine is replaced by an oracle-backed CPU engine (tests may use oracle/), so what is exercised here
is the HOST side of the N>1 path: canonicalise -> partition by owner -> counts all-to-all -> key all-to-all ->
local op -> reverse all-to-all -> un-permute, including chunked overlapped insert, empty ranks and the
collective-with-zero-keys rule (include/kmerhash/distributed_batched_robinhood_map.hpp:843,911-917).
Expected per-rank contents = { (k, c) in the global count map : ref_rank(k, p) == r }.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


class OracleEngine:
    """same interface as distributed.CudaEngine, CPU tensors, oracle arithmetic"""

    def __init__(self, key_words, dist_hash, reducer, canonical_k):
        import oracle
        self.o = oracle.port
        self.device = torch.device("cpu")
        self.W, self.dist_hash, self.canonical_k, self.reducer = key_words, dist_hash, canonical_k, reducer
        self.table = self.o.Map(key_words, reducer)

    def _np(self, t):
        return t.numpy().view(np.uint64)

    def _t(self, a):
        return torch.from_numpy(np.ascontiguousarray(a).view(np.int64))

    def as_keys(self, keys):
        if isinstance(keys, np.ndarray):
            keys = self._t(np.ascontiguousarray(keys, dtype=np.uint64))
        return keys

    def canonicalize(self, keys):
        if not self.canonical_k or keys.shape[0] == 0:
            return keys
        return self._t(self.o.canonical(self.canonical_k, self._np(keys)))

    def ordered(self):
        return self.reducer != "COUNT"

    def as_vals(self, vals):
        return torch.from_numpy(np.ascontiguousarray(vals, dtype=np.uint32).view(np.int32)) if isinstance(vals, np.ndarray) else vals

    def insert_pairs(self, keys, vals, default):
        if not keys.shape[0]:
            return 0
        return self.table.insert(self._np(keys), None if vals is None else vals.numpy().view(np.uint32), default=default)

    def partition(self, keys, p, want_perm, stable=False):
        n = keys.shape[0]
        if n == 0:
            return keys, torch.empty(0, dtype=torch.int32), [0] * p
        r = self.o.rank(self.dist_hash, self._np(keys), p)
        perm = np.argsort(r, kind="stable")
        counts = np.bincount(r, minlength=p).tolist()
        return keys[torch.from_numpy(perm)], torch.from_numpy(perm.astype(np.int32)), counts

    def unpermute(self, vals, perm):
        out = torch.empty_like(vals)
        out[perm.long()] = vals
        return out

    def empty_keys(self, n):
        return torch.empty((n,) if self.W == 1 else (n, self.W), dtype=torch.int64)

    def insert(self, keys, default):
        return self.table.insert(self._np(keys), default=default) if keys.shape[0] else 0

    def count(self, keys):
        if not keys.shape[0]:
            return torch.empty(0, dtype=torch.uint8)
        return torch.from_numpy(np.asarray(self.table.count(self._np(keys)), dtype=np.uint8))

    def find(self, keys, nonexistent):
        if not keys.shape[0]:
            return torch.empty(0, dtype=torch.int32)
        return torch.from_numpy(np.asarray(self.table.find(self._np(keys), nonexistent), dtype=np.uint32).view(np.int32))

    def erase(self, keys):
        return self.table.erase(self._np(keys)) if keys.shape[0] else 0

    def local_size(self):
        return self.table.size()

    def to_vector(self):
        return self.table.to_vector()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_keys(rank, world, W, n):
    rng = np.random.default_rng(100 + rank)
    shape = (n,) if W == 1 else (n, W)
    uniq = rng.integers(0, 2**62, size=shape, dtype=np.uint64)
    if W == 2:
        uniq[:, 1] &= np.uint64((1 << 60) - 1)
    # half of every rank's keys are shared with the other ranks (same seed) so counts add up across ranks
    shared = np.random.default_rng(7).integers(0, 2**62, size=shape, dtype=np.uint64)
    if W == 2:
        shared[:, 1] &= np.uint64((1 << 60) - 1)
    keys = np.concatenate([uniq, shared, uniq[: n // 3]])
    return np.ascontiguousarray(keys[rng.permutation(len(keys))])


def _worker(rank, world, port, W, canonical_k, chunks, empty_rank, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from kmerhash_b200.distributed import DistributedMap
        o = oracle.port
        n = 0 if rank == empty_rank else 600
        keys = _rank_keys(rank, world, W, n) if n else np.zeros((0,) if W == 1 else (0, W), dtype=np.uint64)
        eng = OracleEngine(W, "MURMUR64avx", "COUNT", canonical_k)
        dm = DistributedMap(W, "MURMUR64avx", "MURMUR64avx", "COUNT", canonical_k, engine=eng, chunks=chunks)
        dm.insert(keys.copy())
        # expected: global count map restricted to this rank
        allk = [_rank_keys(r, world, W, 600) for r in range(world) if r != empty_rank]
        allk = np.concatenate(allk)
        if canonical_k:
            allk = o.canonical(canonical_k, allk)
        g = o.Map(W, "COUNT"); g.insert(allk, default=1)
        gk, gv = g.to_vector()
        mine = o.rank("MURMUR64avx", gk, world) == rank
        ek, ev = gk[mine], gv[mine]
        lk, lv = dm.to_vector()
        def srt(k, v):
            order = np.lexsort(k.T[::-1]) if k.ndim == 2 else np.argsort(k)
            return k[order], np.asarray(v)[order]
        ek, ev = srt(ek, ev); lk, lv = srt(np.asarray(lk), lv)
        assert np.array_equal(ek, lk) and np.array_equal(ev, lv), "rank %d partition contents differ" % rank
        assert dm.size() == len(gk)
        # queries come back in INPUT order with global answers
        if len(keys):
            qk = keys[:200].copy()
            qk[::7] ^= np.uint64(1) << np.uint64(61)          # some absent keys
            f = dm.find(qk, 0).numpy().view(np.uint32)
            c = dm.count(qk).numpy()
            qc = o.canonical(canonical_k, qk) if canonical_k else qk
            assert np.array_equal(f, g.find(qc, 0)) and np.array_equal(c, g.count(qc))
        else:
            dm.find(keys, 0); dm.count(keys)                    # collective: every rank calls
        er = dm.erase(keys[:100].copy() if len(keys) else keys)
        tot = torch.tensor([er]); dist.all_reduce(tot)
        ge = g.erase(np.concatenate([(o.canonical(canonical_k, _rank_keys(r, world, W, 600)[:100]) if canonical_k
                                      else _rank_keys(r, world, W, 600)[:100]) for r in range(world) if r != empty_rank]))
        assert int(tot.item()) == ge and dm.size() == g.size()
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "FAIL: %s\n%s" % (e, traceback.format_exc())))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,W,canonical_k,chunks,empty_rank", [
    (2, 1, 0, 1, -1),
    (2, 1, 31, 3, -1),     # canonicalised, chunked overlapped insert
    (2, 2, 63, 1, -1),     # 128-bit keys
    (2, 1, 0, 2, 1),       # one rank brings no keys
    (3, 1, 0, 1, -1),      # non power of two: % p
])
def test_distributed_map_gloo(world, W, canonical_k, chunks, empty_rank):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, W, canonical_k, chunks, empty_rank, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", "rank %d: %s" % (rank, msg)


def _ordered_worker(rank, world, port, reducer, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from kmerhash_b200.distributed import DistributedMap
        o = oracle.port
        batches = []
        for r in range(world):                       # heavy overlap between the ranks: the winner depends on the global order
            rng = np.random.default_rng(40 + r)
            pool = np.random.default_rng(9).integers(0, 2**62, size=300, dtype=np.uint64)
            k = np.ascontiguousarray(pool[rng.integers(0, 300, size=900)])
            v = (np.arange(900, dtype=np.uint32) + np.uint32(100000 * (r + 1)))
            batches.append((k, v))
        eng = OracleEngine(1, "MURMUR64avx", reducer, 0)
        dm = DistributedMap(1, "MURMUR64avx", "MURMUR64avx", reducer, 0, engine=eng)
        for rep in range(2):                         # a second batch: FIRST keeps the old values, LAST takes the new ones
            k, v = batches[rank]
            dm.insert(k.copy(), vals=v + np.uint32(rep))
        # the reference's order at an owner: sender by sender in rank order, input order inside a sender; batch by batch
        g = o.Map(1, reducer)
        for rep in range(2):
            g.insert(np.concatenate([b[0] for b in batches]), np.concatenate([b[1] + np.uint32(rep) for b in batches]))
        gk, gv = g.to_vector()
        mine = o.rank("MURMUR64avx", gk, world) == rank
        lk, lv = dm.to_vector()
        a, b = np.argsort(gk[mine]), np.argsort(np.asarray(lk))
        assert np.array_equal(gk[mine][a], np.asarray(lk)[b]) and np.array_equal(np.asarray(gv)[mine][a], np.asarray(lv)[b])
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        import traceback
        q.put((rank, "FAIL: %s\n%s" % (e, traceback.format_exc())))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,reducer", [(2, "FIRST"), (2, "LAST"), (3, "FIRST")])
def test_distributed_first_last_gloo(world, reducer):
    """(key, value) pairs through the exchange under DiscardReducer / ReplaceReducer: the value a key keeps follows the
    reference's global order (sender rank, then input position), benchmark/BenchmarkDistHashTables.cpp:976-1020"""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_ordered_worker, args=(r, world, port, reducer, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in res:
        assert msg == "ok", "rank %d: %s" % (rank, msg)
