"""One-process-per-GPU distributed map over torch.distributed (NCCL over NVLink / NVSwitch).

Mirror of the reference's distributed container (paths relative to the reference tree):
  dsc::counting_batched_robinhood_map / batched_robinhood_map
      include/kmerhash/distributed_batched_robinhood_map.hpp:748-832 (ctor, sizes), insert :1187 / :2935,
      count :1535, find :1901, erase :2428; insert_p :910-1179, count_p :1258-1516
  exchange: mxx::all2all(counts) :1024 + khmxx::distribute_permuted (include/kmerhash/incremental_mxx.hpp:2998)
            or the overlapped ialltoallv_and_modify (:3437) when chunks > 1
Every rank calls every method (collective semantics, also with 0 keys, :843,911-917).
owner = DistHash_{9876543}(canonical key) & (p-1) or % p -- the reference's own rule, so each rank's
table holds what the reference's rank of the same index holds.  Unlike the reference, query results
come back in INPUT order (the permutation is undone on the source rank).

The compute engine is the CUDA library (kmerhash_b200.ops / HashMap on torch CUDA tensors).  The
`engine` argument exists so the host-side exchange logic can be exercised with gloo on CPU tensors in
the test-suite; the product never passes one.
"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

_TRACE = os.environ.get("KH_TRACE", "0") not in ("", "0")


STAGE_LOG = None      # a list: host wall time (name, ms) of every stage, WITHOUT extra synchronisation (bench.py's per-step breakdown)


class _Stage:
    """KH_TRACE=1: wall time of one stage of a distributed call on rank 0 (device synchronised at both ends).
    STAGE_LOG set: host time of the stage as it runs (every stage ends in a call that synchronises its own stream)."""

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if _TRACE:
            torch.cuda.synchronize()
        if _TRACE or STAGE_LOG is not None:
            self.t0 = time.perf_counter()

    def __exit__(self, *a):
        if _TRACE:
            torch.cuda.synchronize()
            if dist.get_rank() == 0:
                sys.stderr.write("[kh-trace] dist.%-14s %8.3f ms\n" % (self.name, 1e3 * (time.perf_counter() - self.t0)))
        if STAGE_LOG is not None:
            STAGE_LOG.append((self.name, round(1e3 * (time.perf_counter() - self.t0), 3)))


class PeerBuffers:
    """this rank's receive buffer + the other ranks' buffers mapped through CUDA IPC (csrc/exchange.cu: khx_peer_*).
    Collective: every rank of the group constructs it at the same time."""

    def __init__(self, device, rank, p, capacity, key_words, group, query=False):
        import ctypes as C
        from ._lib import lib, check
        self.lib, self.p, self.rank, self.group = lib, p, rank, group
        self.epoch = 0                    # barriers passed on this buffer set (khx_peer_sync_dev)
        self.h = (lib.khx_peer_create_q if query else lib.khx_peer_create)(device, rank, p, int(capacity), key_words)
        ok = 1 if self.h else 0
        dev = torch.device("cuda", device)
        handle = (C.c_ubyte * 64)()
        if ok and lib.khx_peer_handle(self.h, handle) != 0:
            ok = 0
        mine = torch.tensor(list(handle), dtype=torch.uint8, device=dev)
        allh = torch.empty((p, 64), dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(allh, mine, group=group)
        allh = allh.cpu().numpy()
        if ok:
            for r in range(p):
                if r != rank:
                    buf = (C.c_ubyte * 64)(*allh[r].tolist())
                    if lib.khx_peer_open(self.h, r, buf) != 0:
                        ok = 0
                        break
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        self.ok = bool(flag.item())
        self.capacity = lib.khx_peer_capacity(self.h) if self.h else 0

    def recv_ptr(self):
        return self.lib.khx_peer_recv_ptr(self.h)

    def close(self):
        """COLLECTIVE teardown: every rank unmaps its imports, a barrier, then every rank frees its own buffer (freeing an
        exported allocation that a peer still has open is undefined in CUDA)."""
        h, self.h = getattr(self, "h", None), None
        if h:
            self.lib.khx_peer_close_imports(h)
            dist.barrier(group=self.group)
            self.lib.khx_peer_destroy(h)

    def __del__(self):
        # not collective (interpreter exit, error paths): unmap the imports first, then free; a map that was closed
        # explicitly has nothing left to do here
        h, self.h = getattr(self, "h", None), None
        if h:
            try:
                self.lib.khx_peer_close_imports(h)
                self.lib.khx_peer_destroy(h)
            except Exception:
                pass


class CudaEngine:
    """partition / canonicalise / table ops on this rank's GPU through the C ABI"""

    def __init__(self, device, key_words, store_hash, dist_hash, reducer, canonical_k):
        from . import HashMap
        self.device = torch.device("cuda", device)
        self.W, self.dist_hash, self.canonical_k = key_words, dist_hash, canonical_k
        self.table = HashMap(key_words, store_hash, reducer, device=device)

    def as_keys(self, keys):
        if isinstance(keys, np.ndarray):
            keys = torch.from_numpy(np.ascontiguousarray(keys, dtype=np.uint64).view(np.int64))
        return keys.to(self.device, non_blocking=False)

    def canonicalize(self, keys):
        from . import ops
        return ops.canonicalize(keys, self.canonical_k) if (self.canonical_k and keys.shape[0]) else keys

    def partition(self, keys, p, want_perm, stable=False):
        from . import ops
        if keys.shape[0] == 0:
            return keys, torch.empty(0, dtype=torch.int32, device=self.device), [0] * p
        # order inside a destination is irrelevant for COUNT (it commutes) and for queries (they carry their permutation);
        # the order-dependent reducers ask for the stable scatter (permute_by_bucketid, KH/distributed_batched_robinhood_map.hpp:291-380)
        return ops.partition_dev(keys, p, self.dist_hash, want_perm=want_perm, stable=stable)

    def as_vals(self, vals):
        if isinstance(vals, np.ndarray):
            vals = torch.from_numpy(np.ascontiguousarray(vals, dtype=np.uint32).view(np.int32))
        return vals.to(self.device)

    def insert_pairs(self, keys, vals, default):
        return self.table.insert(keys, vals=vals, default=default) if keys.shape[0] else 0

    def ordered(self):
        return self.table.reducer != "COUNT"

    def unpermute(self, vals, perm):
        from . import ops
        return ops.unpermute_dev(vals, perm) if vals.shape[0] else vals

    def empty_keys(self, n):
        shape = (n,) if self.W == 1 else (n, self.W)
        return torch.empty(shape, dtype=torch.int64, device=self.device)

    def empty_like_result(self, n, dtype):
        return torch.empty(n, dtype=dtype, device=self.device)

    def insert(self, keys, default):
        return self.table.insert(keys, default=default) if keys.shape[0] else 0

    # sender / receiver halves with metadata (COUNT tables, p <= 8): see csrc/exchange.cu
    def can_plan(self, p):
        return self.table.reducer == "COUNT" and p <= 8

    def plan_scatter(self, keys, p):
        return self.table.plan_scatter(keys, p, self.dist_hash)

    def insert_recv(self, keys, recv_meta, default):
        return self.table.insert_recv(keys, recv_meta, default) if keys.shape[0] else 0

    def plan(self, keys, p):
        return self.table.plan(keys, p, self.dist_hash)

    def scatter_peers(self, keys, peers, dst_off, dst_cnt):
        self.table.scatter_peers(keys, peers.h, dst_off, dst_cnt, self.dist_hash)

    def insert_recv_ptr(self, ptr, n, recv_meta, default):
        return self.table.insert_recv_ptr(ptr, n, recv_meta, default) if n else 0

    def plan_scatter_slabs(self, keys, peers, slab_cap, parity):
        return self.table.plan_scatter_slabs(keys, peers.h, peers.p, slab_cap, parity, self.dist_hash)

    def insert_recv_slabs(self, peers, slab_cap, parity, counts, recv_meta, default, no_meta=False):
        return self.table.insert_recv_slabs(peers.h, slab_cap, parity, counts, recv_meta, default, no_meta=no_meta) if int(counts.sum()) else 0

    def insert_pipelined(self, keys, peers, slab_cap, parity, chunks, default):
        # KH_PIPE_MODE: "ce" = copy engines move the runs (khx_insert_ce_dev), "sm" = the sender kernels store over NVLink themselves
        ce = os.environ.get("KH_PIPE_MODE", "ce") == "ce"
        got, C, over = self.table.insert_pipelined(keys, peers.h, peers.p, slab_cap, parity, chunks, peers.epoch + 1, default, self.dist_hash,
                                                   copy_engine=ce)
        peers.epoch += chunks
        return got, C, over

    def kmers_scatter_slabs(self, bases, offsets, k, canonical, peers, slab_cap, parity, counts):
        return self.table.kmers_scatter_slabs(bases, offsets, k, canonical, peers.h, slab_cap, parity, counts, self.dist_hash)

    def scatter_slabs(self, keys, peers, slab_cap, parity, counts, reset=True, synced=False):
        self.table.scatter_slabs(keys, peers.h, slab_cap, parity, counts, reset, self.dist_hash, synced=synced)

    def insert_ptr(self, ptr, n, default):
        return self.table.insert_ptr(ptr, n, default) if n else 0

    def push_meta_sync(self, peers, parity, meta):
        peers.epoch += 1
        return self.table.push_meta_sync(peers.h, parity, meta, peers.epoch, peers.p)

    def query_scatter_slabs(self, keys, peers, slab_cap, perm, counts):
        self.table.query_scatter_slabs(keys, peers.h, slab_cap, perm, counts, self.dist_hash)

    def peer_sync(self, peers, counts_row=None):
        peers.epoch += 1
        return self.table.peer_sync(peers.h, peers.epoch, counts_row, peers.p)

    def query_slabs(self, peers, op, slab_cap, counts, nonexistent):
        return self.table.query_slabs(peers.h, op, slab_cap, counts, nonexistent)

    def unpermute_slabs(self, peers, slab_cap, perm, counts, out):
        self.table.unpermute_slabs(peers.h, slab_cap, perm, counts, out)

    def count(self, keys):
        return self.table.count(keys) if keys.shape[0] else self.empty_like_result(0, torch.uint8)

    def find(self, keys, nonexistent):
        return self.table.find(keys, nonexistent) if keys.shape[0] else self.empty_like_result(0, torch.int32)

    def erase(self, keys):
        return self.table.erase(keys) if keys.shape[0] else 0

    def local_size(self):
        return self.table.size()

    def to_vector(self):
        return self.table.to_vector()


class DistributedMap:
    def __init__(self, key_words=1, store_hash="MURMUR64avx", dist_hash="MURMUR64avx", reducer="COUNT",
                 canonical_k=0, device=None, group=None, engine=None, chunks=1):
        self.group = group
        self.rank = dist.get_rank(group)
        self.p = dist.get_world_size(group)
        if engine is None:
            if device is None:
                device = torch.cuda.current_device()
            engine = CudaEngine(device, key_words, store_hash, dist_hash, reducer, canonical_k)
        self.e = engine
        self.W = key_words
        self.chunks = max(1, int(chunks))
        self.last_exchange_bytes = 0      # bytes this rank sent to other ranks in the last call
        # fused partition + exchange over peer memory (NVLink): on when every rank could map every buffer
        self.peer_stores = os.environ.get("KH_PEER_STORES", "1") not in ("0", "")
        self.peers = None
        # single-pass sender into per-sender slabs of the receive buffers (csrc/exchange.cu: scatter_slabs_kernel)
        self.slabs = os.environ.get("KH_SLABS", "1") not in ("0", "")
        self.slab_cap = 0                 # keys per (owner, sender) slab; the buffers hold 2 halves x p slabs
        self.parity = 0
        self.slab_misses = 0              # consecutive batches that overflowed a slab (skewed owners)
        # exchange overlapped with the local insert (csrc/exchange.cu: khx_insert_ce_dev / khx_insert_pipelined_dev): pieces per
        # batch, 1 = off.  Default: 4 pieces with the copy engines on the links from 4 ranks on -- measured on 8 GPUs 3.44 ms per
        # 100 M keys / GPU against 3.75 ms for the one-piece schedule; on 2 GPUs, where the sender kernel is not link-bound, one
        # piece is faster (3.16 against 3.27 ms)
        self.pipe_chunks = int(os.environ.get("KH_PIPE_CHUNKS", "4" if self.p >= 4 else "1"))
        self.pipe_min_keys = 1 << 22
        # above this many keys per rank and batch (slab capacity) the one-piece schedule runs: it is the one measured at 1 B keys per
        # GPU (config 5); the pieces' staging and record buffers grow with the batch
        self.pipe_max_keys = int(os.environ.get("KH_PIPE_MAX_KEYS", str(1 << 29)))
        self._copy_stream = self._lib_stream = None   # host -> device pipeline of the sender (_send_slabs)
        self._dev_keys = self._C = None
        self._exact_dirty = False         # the last batch went through the exact path: its buffer use overlaps both halves
        # queries over peer slabs with direct answer stores (csrc/exchange.cu: khx_query_*)
        self.qslabs = self.slabs
        self.qpeers = None
        self.q_slab_cap = 0
        self.q_misses = 0

    # ---- the exchange ---------------------------------------------------------------------------
    def _counts_a2a(self, counts):
        send = torch.tensor(counts, dtype=torch.int64, device=self.e.device)
        recv = torch.empty_like(send)
        dist.all_to_all_single(recv, send, group=self.group)          # mxx::all2all(send_counts)
        return [int(c) for c in recv.tolist()]

    def _send_keys(self, keys, want_perm, async_op=False):
        """canonicalise -> partition by owner -> all-to-all.  Returns (recv_keys, perm, send_counts, recv_counts, work)."""
        keys = self.e.canonicalize(keys)
        if self.p == 1:                                                # insert_1 / count_1: no exchange
            return keys, None, [keys.shape[0]], [keys.shape[0]], None
        with _Stage("partition"):
            permuted, perm, send_counts = self.e.partition(keys, self.p, want_perm)
        with _Stage("counts a2a"):
            recv_counts = self._counts_a2a(send_counts)
        recv = self.e.empty_keys(sum(recv_counts))
        elem = 8 * self.W
        self.last_exchange_bytes += elem * (sum(send_counts) - send_counts[self.rank])
        with _Stage("keys a2a"):
            work = dist.all_to_all_single(recv, permuted, output_split_sizes=recv_counts, input_split_sizes=send_counts,
                                          group=self.group, async_op=async_op)
        return recv, perm, send_counts, recv_counts, (work, permuted)

    def _reply(self, res, perm, send_counts, recv_counts):
        """results travel back along the reverse all-to-all and are put into input order"""
        if self.p == 1:
            return res
        back = torch.empty(sum(send_counts), dtype=res.dtype, device=self.e.device)
        dist.all_to_all_single(back, res, output_split_sizes=send_counts, input_split_sizes=recv_counts, group=self.group)
        self.last_exchange_bytes += res.element_size() * (sum(recv_counts) - recv_counts[self.rank])
        return self.e.unpermute(back, perm)

    # ---- batch operations (collective) -------------------------------------------------------------
    MAX_BATCH = 1 << 29        # keys per exchange round: receive slabs, 32-bit slab positions and the bulk insert's record positions

    def _insert_ordered(self, keys, vals, default):
        """(key, value) pairs under FIRST / LAST (DiscardReducer / ReplaceReducer, benchmark/BenchmarkDistHashTables.cpp:976-1020):
        which value a key keeps depends on the order in which the owner sees the pairs.  The reference's owner sees them
        sender by sender in rank order, every sender's pairs in its input order (stable permute_by_bucketid + alltoallv,
        KH/distributed_batched_robinhood_map.hpp:291-380,910-1179); the same order is produced here: stable partition,
        all-to-all of keys and of values (NCCL), one local insert over the received sequence."""
        keys = self.e.canonicalize(self.e.as_keys(keys))
        if vals is not None:
            vals = self.e.as_vals(vals)
        if self.p == 1:
            return self.e.insert_pairs(keys, vals, default)
        permuted, perm, send_counts = self.e.partition(keys, self.p, True, stable=True)
        recv_counts = self._counts_a2a(send_counts)
        recv = self.e.empty_keys(sum(recv_counts))
        dist.all_to_all_single(recv, permuted, output_split_sizes=recv_counts, input_split_sizes=send_counts, group=self.group)
        self.last_exchange_bytes += 8 * self.W * (sum(send_counts) - send_counts[self.rank])
        rv = None
        if vals is not None:
            pv = vals[perm.long()] if vals.shape[0] else vals
            rv = torch.empty(sum(recv_counts), dtype=pv.dtype, device=pv.device)
            dist.all_to_all_single(rv, pv, output_split_sizes=recv_counts, input_split_sizes=send_counts, group=self.group)
            self.last_exchange_bytes += 4 * (sum(send_counts) - send_counts[self.rank])
        return self.e.insert_pairs(recv, rv, default)

    def insert(self, keys, default=1, vals=None):
        """returns the number of new keys on THIS rank.  Collective.  vals: one 32-bit value per key (FIRST / LAST tables).  Batches above MAX_BATCH keys go in slices (every rank
        must then bring the same number of slices, i.e. about the same number of keys, as the benchmark drivers do).
        chunks > 1 (NCCL path only) overlaps the all-to-all of chunk i+1 with the local insert of chunk i (the reference's
        ialltoallv_and_modify schedule)."""
        n = int(keys.shape[0])
        if self.p > 1 and n > self.MAX_BATCH:
            slices = (n + self.MAX_BATCH - 1) // self.MAX_BATCH
            added, xb = 0, 0
            for i in range(slices):
                lo, hi = n * i // slices, n * (i + 1) // slices
                added += self.insert(keys[lo:hi], default, None if vals is None else vals[lo:hi])
                xb += self.last_exchange_bytes
            self.last_exchange_bytes = xb
            return added
        self.last_exchange_bytes = 0
        if vals is not None or getattr(self.e, "ordered", lambda: False)():
            return self._insert_ordered(keys, vals, default)
        if self.p > 1 and self.chunks == 1 and getattr(self.e, "can_plan", lambda p: False)(self.p):
            if self.peer_stores and self.slabs and hasattr(self.e, "scatter_slabs"):
                r = self._insert_slabs(keys, default)                  # takes host keys too: the sender runs behind the PCIe copy
                if r is not None:
                    return r
                return self._insert_exact(self._dev_keys, default, ordered=True)
            keys = self.e.canonicalize(self.e.as_keys(keys))
            if self.peer_stores and hasattr(self.e, "scatter_peers"):
                return self._insert_exact(keys, default, ordered=False)
            return self._insert_planned(keys, default)
        keys = self.e.as_keys(keys)
        n_chunks = self.chunks if self.p > 1 else 1
        bounds = [n * i // n_chunks for i in range(n_chunks + 1)]
        added, pending = 0, None
        for i in range(n_chunks):
            recv, _, _, _, w = self._send_keys(keys[bounds[i]:bounds[i + 1]], False, async_op=n_chunks > 1)
            if pending is not None:
                added += self.e.insert(pending, default)               # overlaps the transfer just issued
            if w is not None and n_chunks > 1:
                w[0].wait()
            pending = recv
        with _Stage("local insert"):
            added += self.e.insert(pending, default)
        return added

    def _insert_planned(self, keys, default):
        """NCCL path (no peer access): assign_count_estimate_permute shape: the sender's metadata pass yields, per
        destination, the key count and the HLL registers; they travel with the keys."""
        with _Stage("plan+scatter"):
            permuted, meta, send_counts = self.e.plan_scatter(keys, self.p)
        with _Stage("meta a2a"):
            recv_meta = torch.empty_like(meta)
            dist.all_to_all_single(recv_meta, meta, group=self.group)
            recv_counts = [int(c) for c in recv_meta[:, 0].tolist()]
        recv = self.e.empty_keys(sum(recv_counts))
        self.last_exchange_bytes += 8 * self.W * (sum(send_counts) - send_counts[self.rank]) + 8 * meta.shape[1] * (self.p - 1)
        with _Stage("keys a2a"):
            dist.all_to_all_single(recv, permuted, output_split_sizes=recv_counts, input_split_sizes=send_counts, group=self.group)
        with _Stage("local insert"):
            return self.e.insert_recv(recv, recv_meta, default)

    def _ensure_peers(self, capacity):
        """(re)allocate the receive buffers; collective.  False when IPC / peer access is unavailable."""
        if self.peers is not None:
            self.peers.close()
        self.peers = PeerBuffers(self.e.device.index, self.rank, self.p, int(capacity), self.W, self.group)
        if not self.peers.ok:
            self.peer_stores = False
            self.peers.close(); self.peers = None
            return False
        return True

    HOST_CHUNK = 1 << 23       # keys per piece of the host -> device pipeline

    def _send_slabs(self, keys, parity, counts):
        """keys -> the owners' slabs (khx_scatter_slabs_dev).  Host keys (pinned CPU tensor / numpy array) are copied in pieces
        on a copy stream and the sender kernel of piece i runs behind the copy of piece i+1.  Returns the device keys."""
        dev = self.e.device
        if torch.is_tensor(keys) and keys.is_cuda:
            keys = self.e.canonicalize(keys)
            self.e.scatter_slabs(keys, self.peers, self.slab_cap, parity, counts, reset=True)
            return keys
        if isinstance(keys, np.ndarray):
            keys = torch.from_numpy(np.ascontiguousarray(keys, dtype=np.uint64).view(np.int64))
        n = int(keys.shape[0])
        if self.e.canonical_k or n == 0:
            d = self.e.canonicalize(keys.to(dev))
            self.e.scatter_slabs(d, self.peers, self.slab_cap, parity, counts, reset=True)
            return d
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._lib_stream = torch.cuda.ExternalStream(self.e.table.stream_ptr(), device=dev)
        d = torch.empty(keys.shape, dtype=torch.int64, device=dev)
        torch.cuda.current_stream(dev).synchronize()
        for off in range(0, n, self.HOST_CHUNK):
            m = min(self.HOST_CHUNK, n - off)
            with torch.cuda.stream(self._copy_stream):
                d[off:off + m].copy_(keys[off:off + m], non_blocking=True)
                ev = torch.cuda.Event(); ev.record(self._copy_stream)
            self._lib_stream.wait_event(ev)
            self.e.scatter_slabs(d[off:off + m], self.peers, self.slab_cap, parity, counts, reset=(off == 0), synced=True)
        return d

    def insert_reads(self, bases, offsets, k, canonical=True, default=1):
        """BenchmarkKmerCounter's hot path (KmerParser -> transform_input -> insert, KI/index/kmer_index.hpp:99-225) on device
        buffers: `bases` uint8 CUDA tensor (reads back to back), `offsets` int64 CUDA tensor [n_reads + 1].  With peer
        access the k-mers are generated INSIDE the sender kernel and go straight to their owners' slabs
        (khx_kmers_scatter_slabs_dev); otherwise they are generated first and inserted like any batch.  Collective."""
        from . import ops
        self.last_exchange_bytes = 0
        fused = (self.p > 1 and self.peer_stores and self.slabs and hasattr(self.e, "kmers_scatter_slabs")
                 and getattr(self.e, "can_plan", lambda p: False)(self.p) and not self.e.ordered())
        if not fused:
            return self.insert(ops.kmers_from_reads_dev(bases, offsets, k, canonical), default)
        reads = (bases, offsets, int(k), bool(canonical))
        r = self._insert_slabs(None, default, reads=reads)
        if r is not None:
            return r
        if self._dev_keys is None:                                          # overflow: the exact path wants the k-mers as an array
            self._dev_keys = ops.kmers_from_reads_dev(bases, offsets, k, canonical)
        return self._insert_exact(self._dev_keys, default, ordered=True)

    def _insert_slabs(self, keys, default, reads=None):
        """ONE pass over the keys on the sender: dist hash -> owner -> store into this rank's slab of the owner's buffer
        (NVLink); the flag barrier that follows delivers the p x p count matrix; the receiver's bulk insert reads its p slabs
        in place, its level-1 pass computing the HyperLogLog registers the reference ships with the counts.  Returns None
        when a slab overflowed (every rank sees the same matrix, so every rank takes the exact path for this batch)."""
        p, rank, dev = self.p, self.rank, self.e.device
        n = int(keys.shape[0]) if reads is None else int(reads[0].shape[0])  # reads: every k-mer ends at its own byte
        if self.peers is None or self.slab_cap == 0:
            nmax = torch.tensor([n], dtype=torch.int64, device=dev)
            dist.all_reduce(nmax, op=dist.ReduceOp.MAX, group=self.group)
            self.slab_cap = (int(int(nmax.item()) / p * 1.25) + 4096 + 63) // 64 * 64     # divisible by 2 * chunks (pipelined pieces)
            if not self._ensure_peers(2 * p * self.slab_cap):
                keys = self.e.canonicalize(self.e.as_keys(keys))
                return self._insert_planned(keys, default)
            self.parity, self._exact_dirty = 0, False
        if self._exact_dirty:                                              # a rank may still read the exact-path buffer
            dist.barrier(group=self.group)
            self._exact_dirty = False
        parity = self.parity
        if (reads is None and self.pipe_chunks > 1 and torch.is_tensor(keys) and keys.is_cuda and p <= 8 and hasattr(self.e, "insert_pipelined")
                and self.slab_cap % (2 * self.pipe_chunks) == 0 and self.pipe_min_keys <= p * self.slab_cap <= min(self.pipe_max_keys, 3 << 29)):
            # every rank takes this branch or none: the decision rests on the slab size, which all ranks share, and on where the
            # keys live, which a collective caller keeps uniform (device tensors on every rank, or host arrays on every rank)
            keys = self.e.canonicalize(keys)
            with _Stage("pipelined insert"):
                got, C, over = self.e.insert_pipelined(keys, self.peers, self.slab_cap, parity, self.pipe_chunks, default)
            self.parity ^= 1
            if over:                                                       # same verdict on every rank
                self.slab_misses += 1
                self._C, self._dev_keys = C, keys
                return None
            self.slab_misses = 0
            self.last_exchange_bytes += 8 * self.W * int(C[rank].sum() - C[rank, rank])
            return got
        counts = torch.zeros(p, dtype=torch.int64, device=dev)
        with _Stage("scatter->slabs"):
            if reads is None:
                self._dev_keys = self._send_slabs(keys, parity, counts)
            else:
                self._dev_keys = None
                if not self.e.kmers_scatter_slabs(reads[0], reads[1], reads[2], reads[3], self.peers, self.slab_cap, parity, counts):
                    from . import ops                                      # EOL-dense input: generate, then send
                    self._dev_keys = ops.kmers_from_reads_dev(*reads)
                    self.e.scatter_slabs(self._dev_keys, self.peers, self.slab_cap, parity, counts, reset=True)
        with _Stage("sync+counts"):
            # the flag barrier: every rank's stores have landed and every rank holds the count matrix C[g][r] = keys g sends to r
            C = self.e.peer_sync(self.peers, counts)
        if int(C.max()) > self.slab_cap:                                   # same decision on every rank
            self.slab_misses += 1
            self._C = C
            return None
        self._dev_keys = None
        self.last_batch_keys = int(C[rank].sum())                          # keys (k-mers) this rank contributed
        self.slab_misses = 0
        self.parity ^= 1
        self.last_exchange_bytes += 8 * self.W * int(C[rank].sum() - C[rank, rank])
        with _Stage("local insert"):
            return self.e.insert_recv_slabs(self.peers, self.slab_cap, parity, C[:, rank], None, default, no_meta=True)

    def _insert_exact(self, keys, default, ordered):
        """exact offsets: count matrix -> every key to its final place in the owner's buffer (khx_scatter_peers_dev) -> barrier ->
        insert.  ordered=True: called after a slab overflow, the count matrix of this batch is already known."""
        p, rank, dev = self.p, self.rank, self.e.device
        if ordered:
            C = self._C
            self._dev_keys = None
            need = int(C.sum(axis=0).max())
            grow = need > self.peers.capacity
            if self.slab_misses >= 2:                        # owners stay skewed: keep to the exact path from now on
                self.slabs = False
            elif int(C.max()) * p <= 2 * int(C.sum(axis=1).max()):   # balanced owners, larger batches: larger slabs
                self.slab_cap = (int(int(C.max()) * 1.25) + 4096 + 63) // 64 * 64
                grow = grow or 2 * p * self.slab_cap > self.peers.capacity
            if grow and not self._ensure_peers(max(int(need * 1.25) + 1024, 2 * p * self.slab_cap)):
                return self._insert_planned(keys, default)
            if not grow:
                dist.barrier(group=self.group)               # every rank is past its reads of the previous batch's slabs
        else:
            with _Stage("plan"):
                meta = self.e.plan(keys, p)
            with _Stage("meta gather"):
                allmeta = torch.empty((p,) + tuple(meta.shape), dtype=meta.dtype, device=dev)
                dist.all_gather_into_tensor(allmeta, meta, group=self.group)   # also orders ranks: previous reads are done
                C = allmeta[:, :, 0].cpu().numpy().astype(np.uint64)           # C[g][r] = keys g sends to r
            need = int(C.sum(axis=0).max())
            if self.peers is None or self.peers.capacity < need:               # same decision on every rank
                if not self._ensure_peers(int(need * 1.25) + 1024):            # no IPC / no peer access: NCCL path
                    return self._insert_planned(keys, default)
        self._exact_dirty = True
        self.parity = 0
        dst_off = C[:rank, :].sum(axis=0)
        dst_cnt = C[rank, :]
        self.last_exchange_bytes += 8 * self.W * int(dst_cnt.sum() - dst_cnt[rank])
        with _Stage("scatter->peers"):
            self.e.scatter_peers(keys, self.peers, dst_off, dst_cnt)
            dist.barrier(group=self.group)                                     # every rank's stores have landed
        with _Stage("local insert"):
            return self.e.insert_ptr(self.peers.recv_ptr(), int(C[:, rank].sum()), default)

    def close(self):
        """COLLECTIVE: releases the IPC-mapped exchange buffers (unmap imports -> barrier -> free)"""
        for name in ("peers", "qpeers"):
            pb = getattr(self, name, None)
            setattr(self, name, None)
            if pb is not None:
                pb.close()

    def _query_slabs(self, op, keys, nonexistent=0):
        """find / count / erase without NCCL on the data path: keys go to the owners by peer stores (one kernel), every
        owner's lookup kernel stores its answers straight into the asker's reply region, the asker un-permutes.  The two
        barriers (behind the key stores, with the p x p counts; behind the answers) are flag kernels over peer memory.
        Returns None when this op has to take the NCCL path (no peer access, a slab overflowed)."""
        self.last_query_path = "nccl"
        if not (self.p > 1 and self.peer_stores and self.qslabs and hasattr(self.e, "query_scatter_slabs")):
            return None
        p, rank, dev = self.p, self.rank, self.e.device
        keys = self.e.canonicalize(keys)
        n = int(keys.shape[0])
        if self.qpeers is None:
            nmax = torch.tensor([n], dtype=torch.int64, device=dev)
            dist.all_reduce(nmax, op=dist.ReduceOp.MAX, group=self.group)
            self.q_slab_cap = int(int(nmax.item()) / p * 1.25) + 4096
            self.qpeers = PeerBuffers(dev.index, rank, p, p * self.q_slab_cap, self.W, self.group, query=True)
            if not self.qpeers.ok:
                self.qslabs = False
                self.qpeers.close(); self.qpeers = None
                return None
        cap = self.q_slab_cap
        with _Stage("q.alloc"):
            perm = torch.empty(p * cap, dtype=torch.int32, device=dev)
            counts = torch.zeros(p, dtype=torch.int64, device=dev)
        with _Stage("q.scatter"):
            self.e.query_scatter_slabs(keys, self.qpeers, cap, perm, counts)
        with _Stage("q.sync+counts"):
            C = self.e.peer_sync(self.qpeers, counts)                         # every rank's key stores have landed; C[g][r]
        if int(C.max()) > cap:                                                # same decision on every rank
            if int(C.max()) * p <= 2 * int(C.sum(axis=1).max()):              # balanced owners, larger batches: larger slabs
                self.q_slab_cap = int(int(C.max()) * 1.25) + 4096
                self.qpeers.close()
                self.qpeers = PeerBuffers(dev.index, rank, p, p * self.q_slab_cap, self.W, self.group, query=True)
                if not self.qpeers.ok:
                    self.qslabs = False
                    self.qpeers.close(); self.qpeers = None
            else:
                self.q_misses += 1
                if self.q_misses >= 2:                                        # owners stay skewed: NCCL path from now on
                    self.qslabs = False
            return None
        self.q_misses = 0
        self.last_query_path = "slab"
        elem = {"find": 4, "count": 1, "erase": 0}[op]
        self.last_exchange_bytes += (8 * self.W + elem) * int(C[rank].sum() - C[rank, rank])
        with _Stage("q.lookup"):
            hits = self.e.query_slabs(self.qpeers, op, cap, C[:, rank], nonexistent)
        with _Stage("q.sync"):
            self.e.peer_sync(self.qpeers)                                     # every owner's answers have landed
        if op == "erase":
            return hits
        out = torch.empty(n, dtype=torch.int32 if op == "find" else torch.uint8, device=dev)
        with _Stage("q.unpermute"):
            self.e.unpermute_slabs(self.qpeers, cap, perm, counts, out)
        return out

    def count(self, keys):
        keys = self.e.as_keys(keys)
        self.last_exchange_bytes = 0
        r = self._query_slabs("count", keys)
        if r is not None:
            return r
        recv, perm, sc, rc, _ = self._send_keys(keys, True)
        return self._reply(self.e.count(recv), perm, sc, rc)

    def find(self, keys, nonexistent=0):
        keys = self.e.as_keys(keys)
        self.last_exchange_bytes = 0
        r = self._query_slabs("find", keys, nonexistent)
        if r is not None:
            return r
        recv, perm, sc, rc, _ = self._send_keys(keys, True)
        return self._reply(self.e.find(recv, nonexistent), perm, sc, rc)

    def erase(self, keys):
        keys = self.e.as_keys(keys)
        self.last_exchange_bytes = 0
        r = self._query_slabs("erase", keys)
        if r is not None:
            return r
        recv, _, _, _, _ = self._send_keys(keys, False)
        return self.e.erase(recv)

    # ---- sizes ------------------------------------------------------------------------------------------
    def local_size(self):
        return self.e.local_size()

    def size(self):
        t = torch.tensor([self.local_size()], dtype=torch.int64, device=self.e.device)
        dist.all_reduce(t, group=self.group)
        return int(t.item())

    def get_local_container(self):
        return getattr(self.e, "table", None)

    def to_vector(self):
        """this rank's (keys, vals)"""
        return self.e.to_vector()
