"""Host-side mirror of the reference's container interface over the C ABI.

Class / method names follow the reference (paths relative to the reference tree):
  hashmap_robinhood_offsets_count / hashmap_robinhood_offsets / ..._reduction
      include/kmerhash/robinhood_offset_hashmap_ptr.hpp:196-202,4094,4099
  hashmap_robinhood_doubling      include/kmerhash/hashmap_robinhood.hpp:126
  hashmap_linearprobe_doubling    include/kmerhash/hashmap_linearprobe.hpp:98
Batch methods take either numpy arrays (HOST memory; the library moves the data) or torch CUDA
tensors (DEVICE memory on the table's GPU; zero copy).  Keys are uint64 words: shape (n,) for
k <= 32, (n, 2) for k <= 64 -- the memory layout of bliss::common::Kmer<K, DNA, uint64_t>.
"""
import ctypes as C

import numpy as np

from ._lib import lib, check, HASH_KINDS, REDUCERS, MAP_KINDS, DEFAULT_SEED, DEFAULT_CRC_SEED

try:  # torch is plumbing only (device buffers); the table itself never needs it
    import torch
except Exception:  # pragma: no cover
    torch = None


def _is_torch(x):
    return torch is not None and isinstance(x, torch.Tensor)


def _host_keys(keys, W):
    a = np.ascontiguousarray(keys, dtype=np.uint64)
    if W == 1:
        if a.ndim == 2 and a.shape[1] == 1:
            a = a[:, 0].copy()
        if a.ndim != 1:
            raise ValueError("expected keys of shape (n,) for 1-word keys")
    else:
        if a.ndim != 2 or a.shape[1] != W:
            raise ValueError("expected keys of shape (n, %d)" % W)
    return a, a.shape[0]


def _dev_keys(keys, W, device):
    if not keys.is_cuda or keys.device.index != device:
        raise ValueError("device keys must live on cuda:%d" % device)
    if keys.dtype not in (torch.int64, torch.uint64):
        raise ValueError("device keys must be int64/uint64 words")
    if not keys.is_contiguous():
        keys = keys.contiguous()
    n = keys.shape[0]
    if W == 2 and (keys.dim() != 2 or keys.shape[1] != 2):
        raise ValueError("expected keys of shape (n, 2)")
    if W == 1 and keys.dim() != 1:
        raise ValueError("expected keys of shape (n,)")
    torch.cuda.current_stream(device).synchronize()   # the library works on its own streams
    return keys, n


class HashMap:
    """One table on one GPU.  `reducer` selects COUNT (std::plus), FIRST (Discard) or LAST (Replace)."""

    def __init__(self, key_words=1, hash_kind="MURMUR64avx", reducer="COUNT", map_kind="ROBINHOOD_OFFSETS",
                 device=0, seed=None, min_load=None, max_load=None):
        if seed is None:
            seed = DEFAULT_CRC_SEED if hash_kind == "CRC32C" else DEFAULT_SEED
        self.W, self.device = int(key_words), int(device)
        self.hash_kind, self.reducer, self.map_kind = hash_kind, reducer, map_kind
        self._h = lib.kh_create(self.device, MAP_KINDS[map_kind], self.W, HASH_KINDS[hash_kind], seed, REDUCERS[reducer])
        if not self._h:
            check(-3 if lib.kh_device_count() <= 0 else -1, "kh_create")
        if min_load is not None or max_load is not None:
            self.set_load_factors(min_load if min_load is not None else 0.2, max_load if max_load is not None else 0.8)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib.kh_destroy(h)

    # ---- sizing (reference: set_max/min_load_factor, reserve, rehash, capacity, size, clear) ----
    def set_load_factors(self, min_load, max_load):
        check(lib.kh_set_load_factors(self._h, min_load, max_load), "set_load_factors")

    def set_insert_lookahead(self, n):
        lib.kh_set_lookahead(self._h, int(n), 0)

    def set_query_lookahead(self, n):
        lib.kh_set_lookahead(self._h, 0, int(n))

    def set_geometry(self, lane_bytes, lanes_per_bucket):
        check(lib.kh_set_geometry(self._h, lane_bytes, lanes_per_bucket), "set_geometry")

    def set_insert_window(self, enable=True, window_bytes=32 << 20, min_keys=1 << 22):
        """L2-windowed bulk insert switch / thresholds (results do not depend on it)"""
        check(lib.kh_set_insert_window(self._h, int(enable), int(window_bytes), int(min_keys)), "set_insert_window")

    def stream_ptr(self):
        """cudaStream_t (as int) the device-pointer calls of this table launch on"""
        return lib.kh_stream(self._h)

    def kernel_timing(self, enable=True):
        check(lib.kh_kernel_timing(self._h, int(enable)), "kernel_timing")

    def kernel_time(self, which="insert"):
        """(sum of ms, launches) of the insert / lookup probe kernels since kernel_timing(True)"""
        ms, n = C.c_double(0), C.c_uint64(0)
        check(lib.kh_kernel_time(self._h, {"insert": 0, "lookup": 1, "scatter": 2, "hll": 3, "rebuild": 4, "level2": 5}[which], C.byref(ms), C.byref(n)), "kernel_time")
        return ms.value, n.value

    def reserve(self, n):
        check(lib.kh_reserve(self._h, int(n)), "reserve")

    def rehash(self, slots):
        check(lib.kh_rehash(self._h, int(slots)), "rehash")

    def capacity(self):
        return lib.kh_capacity(self._h)

    def size(self):
        return lib.kh_size(self._h)

    __len__ = size

    def get_load_factor(self):
        return lib.kh_load_factor(self._h)

    def clear(self):
        check(lib.kh_clear(self._h), "clear")

    # ---- estimate (reference: get_hll().update + estimate) ----
    def estimate_distinct(self, keys, return_registers=False):
        regs = np.zeros(4096, dtype=np.uint8)
        if _is_torch(keys):
            k, n = _dev_keys(keys, self.W, self.device)
            est = lib.kh_estimate_distinct_dev(self._h, k.data_ptr(), n, regs.ctypes.data)
        else:
            k, n = _host_keys(keys, self.W)
            est = lib.kh_estimate_distinct_host(self._h, k.ctypes.data, n, regs.ctypes.data)
        if est < 0:
            check(-3, "estimate_distinct")
        return (est, regs) if return_registers else est

    # ---- insert (reference: insert / insert_no_estimate) ----
    def insert(self, keys, vals=None, default=1, estimate=True):
        """Returns the number of keys that were new."""
        if _is_torch(keys):
            k, n = _dev_keys(keys, self.W, self.device)
            vp = None
            if vals is not None:
                vals = vals.contiguous()
                assert vals.is_cuda and vals.dtype in (torch.int32, torch.uint32) and vals.shape[0] == n
                vp = vals.data_ptr()
            return check(lib.kh_insert_dev(self._h, k.data_ptr(), vp, n, default, int(estimate)), "insert")
        k, n = _host_keys(keys, self.W)
        vp = None
        if vals is not None:
            vals = np.ascontiguousarray(vals, dtype=np.uint32)
            assert vals.shape[0] == n
            vp = vals.ctypes.data
        return check(lib.kh_insert_host(self._h, k.ctypes.data, vp, n, default, int(estimate)), "insert")

    # ---- one rank of a multi-process distributed insert (include/kmerhash_b200.h: khx_*) ----
    def plan_scatter(self, keys, p, dist_hash="MURMUR64avx", dist_seed=9876543):
        """sender half: (keys permuted so that every destination's keys are contiguous, metadata tensor
        [p, khx_meta_words] int64 on the device, counts[p] python list).  CUDA tensors only."""
        k, n = _dev_keys(keys, self.W, self.device)
        out = torch.empty_like(k)
        meta = torch.empty((p, lib.khx_meta_words()), dtype=torch.int64, device=k.device)
        counts = np.zeros(p, dtype=np.uint64)
        check(lib.khx_plan_scatter_dev(self._h, HASH_KINDS[dist_hash], dist_seed, p, k.data_ptr(), n, meta.data_ptr(),
                                       out.data_ptr(), counts.ctypes.data), "plan_scatter")
        return out, meta, [int(c) for c in counts]

    def plan(self, keys, p, dist_hash="MURMUR64avx", dist_seed=9876543):
        """metadata pass alone: [p, khx_meta_words] int64 device tensor (count, region histogram, HLL registers per destination)"""
        k, n = _dev_keys(keys, self.W, self.device)
        meta = torch.empty((p, lib.khx_meta_words()), dtype=torch.int64, device=k.device)
        check(lib.khx_plan_dev(self._h, HASH_KINDS[dist_hash], dist_seed, p, k.data_ptr(), n, meta.data_ptr()), "plan")
        return meta

    def scatter_peers(self, keys, peer_handle, dst_offsets, dst_counts, dist_hash="MURMUR64avx", dist_seed=9876543):
        """fused partition + exchange: store every key into its owner's receive buffer (peer memory)"""
        k, n = _dev_keys(keys, self.W, self.device)
        off = np.ascontiguousarray(dst_offsets, dtype=np.uint64)
        cnt = np.ascontiguousarray(dst_counts, dtype=np.uint64)
        check(lib.khx_scatter_peers_dev(self._h, HASH_KINDS[dist_hash], dist_seed, k.data_ptr(), n, peer_handle,
                                        off.ctypes.data, cnt.ctypes.data), "scatter_peers")

    def plan_scatter_slabs(self, keys, peer_handle, p, slab_cap, parity, dist_hash="MURMUR64avx", dist_seed=9876543):
        """single-pass sender: metadata + rank scatter into this rank's slab of every owner's receive buffer
        (half `parity`); returns the [p, khx_meta_words] metadata tensor"""
        k, n = _dev_keys(keys, self.W, self.device)
        meta = torch.empty((p, lib.khx_meta_words()), dtype=torch.int64, device=k.device)
        check(lib.khx_plan_scatter_slabs_dev(self._h, HASH_KINDS[dist_hash], dist_seed, k.data_ptr(), n, peer_handle,
                                             int(slab_cap), int(parity), meta.data_ptr()), "plan_scatter_slabs")
        return meta

    def scatter_slabs(self, keys, peer_handle, slab_cap, parity, counts, reset=True, dist_hash="MURMUR64avx", dist_seed=9876543, synced=False):
        """insert sender without metadata (khx_scatter_slabs_dev): keys -> this rank's slab of every owner's receive buffer;
        `counts` (int64 CUDA tensor, p entries) accumulates this rank's row of the count matrix.  Asynchronous on the
        table's stream; synced=True: the caller already ordered the keys before that stream."""
        if synced:
            k, n = keys, keys.shape[0]
        else:
            k, n = _dev_keys(keys, self.W, self.device)
        assert counts.is_cuda and counts.dtype == torch.int64
        check(lib.khx_scatter_slabs_dev(self._h, HASH_KINDS[dist_hash], dist_seed, k.data_ptr(), n, peer_handle, int(slab_cap),
                                        int(parity), counts.data_ptr(), int(reset)), "scatter_slabs")

    def insert_pipelined(self, keys, peer_handle, p, slab_cap, parity, chunks, epoch0, default=1, dist_hash="MURMUR64avx", dist_seed=9876543,
                         copy_engine=False):
        """one rank of the pipelined distributed insert (khx_insert_pipelined_dev): returns (new keys, p x p count matrix,
        overflowed)"""
        k, n = _dev_keys(keys, self.W, self.device)
        mat = np.zeros((p, p), dtype=np.uint64)
        over = C.c_int(0)
        fn = lib.khx_insert_ce_dev if copy_engine else lib.khx_insert_pipelined_dev
        got = check(fn(self._h, peer_handle, HASH_KINDS[dist_hash], dist_seed, k.data_ptr(), n, int(slab_cap),
                                                 int(parity), int(chunks), int(epoch0), default, mat.ctypes.data, C.byref(over)),
                    "insert_pipelined")
        return got, mat, bool(over.value)

    def kmers_scatter_slabs(self, bases, offsets, k, canonical, peer_handle, slab_cap, parity, counts, dist_hash="MURMUR64avx", dist_seed=9876543):
        """k-mer generation fused with the sender (khx_kmers_scatter_slabs_dev).  bases: uint8 CUDA tensor, offsets: int64 CUDA
        tensor.  Returns False when the input is EOL-dense and the caller has to take the two-step path."""
        assert bases.is_cuda and bases.dtype == torch.uint8 and offsets.is_cuda and offsets.dtype == torch.int64
        torch.cuda.current_stream(self.device).synchronize()
        rc = lib.khx_kmers_scatter_slabs_dev(self._h, bases.data_ptr(), offsets.data_ptr(), offsets.shape[0] - 1, int(k), int(canonical),
                                             HASH_KINDS[dist_hash], dist_seed, peer_handle, int(slab_cap), int(parity), counts.data_ptr())
        if rc == -4:
            return False
        check(rc, "kmers_scatter_slabs")
        return True

    def insert_ptr(self, ptr, n, default=1, estimate=True):
        """insert n keys lying at a raw device pointer (a peer receive buffer)"""
        torch.cuda.current_stream(self.device).synchronize()
        return check(lib.kh_insert_dev(self._h, ptr, None, int(n), default, int(estimate)), "insert")

    def insert_recv_slabs(self, peer_handle, slab_cap, parity, counts, recv_meta=None, default=1, no_meta=False):
        """receiver half reading the p slabs of this rank's receive buffer in place; counts[g] = keys sender g stored.
        recv_meta: [p, khx_meta_words] int64 CUDA tensor, or None = the buffer's own metadata area (khx_push_meta_sync_dev)"""
        cnt = np.ascontiguousarray(counts, dtype=np.uint64)
        if no_meta:
            mp = None
        elif recv_meta is None:
            mp = lib.khx_peer_meta_ptr(peer_handle, int(parity))
        else:
            assert recv_meta.is_cuda and recv_meta.dtype == torch.int64 and recv_meta.is_contiguous()
            assert cnt.shape[0] == recv_meta.shape[0]
            mp = recv_meta.data_ptr()
        torch.cuda.current_stream(self.device).synchronize()
        return check(lib.khx_insert_recv_slabs_dev(self._h, peer_handle, int(slab_cap), int(parity), cnt.ctypes.data, mp, default),
                     "insert_recv_slabs")

    def push_meta_sync(self, peer_handle, parity, meta, epoch, p):
        """metadata blocks to their owners by peer stores + barrier; returns the p x p count matrix [sender][owner]"""
        assert meta.is_cuda and meta.dtype == torch.int64 and meta.is_contiguous()
        mat = np.zeros((p, p), dtype=np.uint64)
        torch.cuda.current_stream(self.device).synchronize()
        check(lib.khx_push_meta_sync_dev(self._h, peer_handle, int(parity), meta.data_ptr(), int(epoch), mat.ctypes.data), "push_meta_sync")
        return mat

    # query exchange over peer slabs (khx_query_*): see distributed.py
    def query_scatter_slabs(self, keys, peer_handle, slab_cap, perm, counts, dist_hash="MURMUR64avx", dist_seed=9876543):
        k, n = _dev_keys(keys, self.W, self.device)
        assert perm.is_cuda and perm.dtype == torch.int32 and counts.is_cuda and counts.dtype == torch.int64
        check(lib.khx_query_scatter_slabs_dev(self._h, HASH_KINDS[dist_hash], dist_seed, k.data_ptr(), n, peer_handle, int(slab_cap),
                                              perm.data_ptr(), counts.data_ptr()), "query_scatter_slabs")

    def peer_sync(self, peer_handle, epoch, counts_row=None, p=0):
        """barrier over peer memory; with counts_row (int64 CUDA tensor, p entries) returns the p x p matrix [sender][owner]"""
        mat = np.zeros((p, p), dtype=np.uint64) if counts_row is not None else None
        check(lib.khx_peer_sync_dev(self._h, peer_handle, int(epoch), counts_row.data_ptr() if counts_row is not None else None,
                                    mat.ctypes.data if mat is not None else None), "peer_sync")
        return mat

    def query_slabs(self, peer_handle, op, slab_cap, counts, nonexistent=0):
        cnt = np.ascontiguousarray(counts, dtype=np.uint64)
        torch.cuda.current_stream(self.device).synchronize()
        return check(lib.khx_query_slabs_dev(self._h, peer_handle, {"find": 0, "count": 1, "erase": 2}[op], int(slab_cap),
                                             cnt.ctypes.data, int(nonexistent) & 0xFFFFFFFF), "query_slabs")

    def unpermute_slabs(self, peer_handle, slab_cap, perm, counts, out):
        check(lib.khx_unpermute_slabs_dev(self._h, peer_handle, out.element_size(), int(slab_cap), perm.data_ptr(), counts.data_ptr(),
                                          out.shape[0], out.data_ptr()), "unpermute_slabs")

    def insert_recv_ptr(self, ptr, n, recv_meta, default=1):
        """receiver half on a raw device pointer (the peer receive buffer)"""
        assert recv_meta.is_cuda and recv_meta.dtype == torch.int64 and recv_meta.is_contiguous()
        torch.cuda.current_stream(self.device).synchronize()
        return check(lib.khx_insert_recv_dev(self._h, ptr, n, recv_meta.data_ptr(), recv_meta.shape[0], default), "insert_recv")

    def insert_recv(self, keys, recv_meta, default=1):
        """receiver half: insert the received keys, presized and region-sorted from the senders' metadata"""
        k, n = _dev_keys(keys, self.W, self.device)
        assert recv_meta.is_cuda and recv_meta.dtype == torch.int64 and recv_meta.is_contiguous()
        return check(lib.khx_insert_recv_dev(self._h, k.data_ptr(), n, recv_meta.data_ptr(), recv_meta.shape[0], default),
                     "insert_recv")

    def insert_no_estimate(self, keys, vals=None, default=1):
        return self.insert(keys, vals, default, estimate=False)

    # ---- queries ----
    def count(self, keys, out=None):
        """0/1 per key (reference: count(out, b, e) / exists).  `out`: optional preallocated (pinned) uint8 array."""
        if _is_torch(keys):
            k, n = _dev_keys(keys, self.W, self.device)
            out = torch.empty(n, dtype=torch.uint8, device=k.device)
            check(lib.kh_count_dev(self._h, k.data_ptr(), n, out.data_ptr()), "count")
            return out
        k, n = _host_keys(keys, self.W)
        if out is None:
            out = np.empty(n, dtype=np.uint8)
        assert out.dtype == np.uint8 and out.shape[0] >= n and out.flags.c_contiguous
        check(lib.kh_count_host(self._h, k.ctypes.data, n, out.ctypes.data), "count")
        return out

    exists = count

    def find(self, keys, nonexistent=0, out=None):
        """value per key, `nonexistent` where absent (reference: find<OT>(out, b, e, nonexistent))."""
        if _is_torch(keys):
            k, n = _dev_keys(keys, self.W, self.device)
            out = torch.empty(n, dtype=torch.int32, device=k.device)
            check(lib.kh_find_dev(self._h, k.data_ptr(), n, out.data_ptr(), nonexistent), "find")
            return out
        k, n = _host_keys(keys, self.W)
        if out is None:
            out = np.empty(n, dtype=np.uint32)
        assert out.dtype == np.uint32 and out.shape[0] >= n and out.flags.c_contiguous
        check(lib.kh_find_host(self._h, k.ctypes.data, n, out.ctypes.data, nonexistent), "find")
        return out

    def find_existing(self, keys):
        """(keys, vals) of the hits only, in query order (reference: find_existing)."""
        k, n = _host_keys(keys, self.W)
        ok = np.empty((max(n, 1), self.W), dtype=np.uint64)
        ov = np.empty(max(n, 1), dtype=np.uint32)
        c = check(lib.kh_find_existing_host(self._h, k.ctypes.data, n, ok.ctypes.data, ov.ctypes.data), "find_existing")
        return (ok[:c] if self.W > 1 else ok[:c, 0].copy()), ov[:c]

    def update(self, keys, vals):
        k, n = _host_keys(keys, self.W)
        vals = np.ascontiguousarray(vals, dtype=np.uint32)
        return check(lib.kh_update_host(self._h, k.ctypes.data, vals.ctypes.data, n), "update")

    def erase(self, keys):
        if _is_torch(keys):
            k, n = _dev_keys(keys, self.W, self.device)
            return check(lib.kh_erase_dev(self._h, k.data_ptr(), n), "erase")
        k, n = _host_keys(keys, self.W)
        return check(lib.kh_erase_host(self._h, k.ctypes.data, n), "erase")

    def to_vector(self):
        """(keys, vals), unordered (reference: to_vector)."""
        n = self.size()
        ok = np.empty((max(n, 1), self.W), dtype=np.uint64)
        ov = np.empty(max(n, 1), dtype=np.uint32)
        c = check(lib.kh_export_host(self._h, ok.ctypes.data, ov.ctypes.data, n), "to_vector")
        c = min(c, n)
        return (ok[:c] if self.W > 1 else ok[:c, 0].copy()), ov[:c]

    def keys(self):
        return self.to_vector()[0]


# ---- the reference's container names --------------------------------------------------------
def hashmap_robinhood_offsets_count(key_words=1, hash_kind="MURMUR64avx", device=0, **kw):
    return HashMap(key_words, hash_kind, "COUNT", "ROBINHOOD_OFFSETS", device, **kw)


def hashmap_robinhood_offsets(key_words=1, hash_kind="MURMUR64avx", device=0, **kw):
    return HashMap(key_words, hash_kind, "FIRST", "ROBINHOOD_OFFSETS", device, **kw)


def hashmap_robinhood_offsets_reduction(key_words=1, hash_kind="MURMUR64avx", reducer="LAST", device=0, **kw):
    return HashMap(key_words, hash_kind, reducer, "ROBINHOOD_OFFSETS", device, **kw)


def hashmap_robinhood_doubling(key_words=1, hash_kind="MURMUR64", device=0, **kw):
    return HashMap(key_words, hash_kind, "FIRST", "ROBINHOOD", device, **kw)


def hashmap_linearprobe_doubling(key_words=1, hash_kind="MURMUR64", device=0, **kw):
    return HashMap(key_words, hash_kind, "FIRST", "LINEARPROBE", device, **kw)
