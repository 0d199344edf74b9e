"""Batch operators around the table: hashing, k-mer generation, canonicalisation, rank
assignment / partitioning, the single-process multi-GPU map and the roofline probes.
Each function names the reference interface it stands in for (see include/kmerhash_b200.h)."""
import numpy as np

from ._lib import lib, check, HASH_KINDS, REDUCERS, DEFAULT_SEED, DEFAULT_CRC_SEED, DIST_SEED
from .hashmap import _host_keys, _is_torch, _dev_keys, torch


def kmer_words(k):
    return (2 * k + 63) // 64


def _shape_keys(keys):
    a = np.ascontiguousarray(keys, dtype=np.uint64)
    W = 1 if a.ndim == 1 else a.shape[1]
    return a, W, a.shape[0]


def hash_batch(keys, hash_kind="MURMUR64avx", seed=None, device=0):
    """fsc::hash::<kind><Kmer>(seed) over a batch (include/kmerhash/hash_new.hpp batch mode); u64 out."""
    if seed is None:
        seed = DEFAULT_CRC_SEED if hash_kind == "CRC32C" else DEFAULT_SEED
    a, W, n = _shape_keys(keys)
    out = np.empty(n, dtype=np.uint64)
    check(lib.kh_hash_host(device, HASH_KINDS[hash_kind], seed, W, a.ctypes.data, n, out.ctypes.data), "hash_batch")
    return out


ALPHABETS = {"DNA": 4, "DNA5": 5, "DNA16": 16, 4: 4, 5: 5, 16: 16}


def kmers_from_reads(bases, offsets, k, canonical=False, device=0, alphabet="DNA"):
    """KmerParser over N-split sequences (+ lex_less): all k-mers of every read, in read order.  alphabet: "DNA" (2 bits, k <= 64),
    "DNA5" (3 bits, k <= 42), "DNA16" (4 bits, k <= 32) -- bliss::common::Kmer<k, Alphabet, uint64_t> of up to two words."""
    if isinstance(bases, str):
        bases = bases.encode()
    buf = np.frombuffer(bases, dtype=np.uint8) if isinstance(bases, (bytes, bytearray)) else np.ascontiguousarray(bases, dtype=np.uint8)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    n_reads = offsets.shape[0] - 1
    al = ALPHABETS[alphabet]
    W = lib.kh_kmer_words_alpha(al, k)
    total = check(lib.kh_kmers_from_reads_alpha_host(device, al, buf.ctypes.data, offsets.ctypes.data, n_reads, k, int(canonical), None, 0),
                  "kmers_from_reads(count)")
    out = np.empty((max(total, 1), W), dtype=np.uint64)
    if total:
        check(lib.kh_kmers_from_reads_alpha_host(device, al, buf.ctypes.data, offsets.ctypes.data, n_reads, k, int(canonical),
                                                 out.ctypes.data, total), "kmers_from_reads")
    return out[:total] if W > 1 else out[:total, 0].copy()


def canonicalize_alpha(kmers, k, alphabet, device=0, revcomp_only=False):
    """lex_less / reverse_complement for DNA5 / DNA16 (and DNA) k-mers on host arrays"""
    al = ALPHABETS[alphabet]
    a = np.ascontiguousarray(kmers, dtype=np.uint64)
    out = np.empty_like(a)
    fn = lib.kh_revcomp_alpha_host if revcomp_only else lib.kh_canonicalize_alpha_host
    check(fn(device, al, k, a.ctypes.data, a.shape[0], out.ctypes.data), "canonicalize_alpha")
    return out


def kmers_from_reads_dev(bases, offsets, k, canonical=False):
    """Device-resident variant: `bases` uint8 CUDA tensor, `offsets` int64 CUDA tensor -> int64 CUDA tensor."""
    dev = bases.device.index
    n_reads = offsets.shape[0] - 1
    W = kmer_words(k)
    torch.cuda.current_stream(dev).synchronize()
    # one pass: every k-mer ends at its own byte, so the byte count bounds the output; the tile kernel keeps the order
    cap = int(bases.shape[0])
    out = torch.empty((cap, W) if W > 1 else (cap,), dtype=torch.int64, device=bases.device)
    total = 0
    if cap and n_reads:
        total = check(lib.kh_kmers_from_reads_dev(dev, bases.data_ptr(), offsets.data_ptr(), n_reads, k, int(canonical),
                                                  out.data_ptr(), cap), "kmers_from_reads")
    return out[:total]


def canonicalize(kmers, k, device=0):
    """lex_less: min(kmer, reverse_complement) (ext/kmerind/src/common/kmer_transform.hpp:109-125)."""
    if _is_torch(kmers):
        a, n = _dev_keys(kmers, kmer_words(k), kmers.device.index)
        out = torch.empty_like(a)
        check(lib.kh_canonicalize_dev(a.device.index, k, a.data_ptr(), n, out.data_ptr()), "canonicalize")
        return out
    a, n = _host_keys(kmers, kmer_words(k))
    out = np.empty_like(a)
    check(lib.kh_canonicalize_host(device, k, a.ctypes.data, n, out.ctypes.data), "canonicalize")
    return out


def revcomp(kmers, k, device=0):
    """Kmer::reverse_complement (ext/kmerind/src/common/kmer.hpp:1078-1087)."""
    a, n = _host_keys(kmers, kmer_words(k))
    out = np.empty_like(a)
    check(lib.kh_revcomp_host(device, k, a.ctypes.data, n, out.ctypes.data), "revcomp")
    return out


def rank_of(keys, p, hash_kind="MURMUR64avx", seed=DIST_SEED, device=0):
    """owner rank per key: DistHash_seed(key) & (p-1) or % p (distributed_batched_robinhood_map.hpp:164-178,652,749)."""
    a, W, n = _shape_keys(keys)
    out = np.empty(n, dtype=np.uint32)
    check(lib.kh_rank_host(device, HASH_KINDS[hash_kind], seed, W, p, a.ctypes.data, n, out.ctypes.data), "rank_of")
    return out


def partition_dev(keys, p, hash_kind="MURMUR64avx", seed=DIST_SEED, want_perm=True, stable=True):
    """Partition of a CUDA key tensor by owner rank (assign_count_permute + permute_by_bucketid).
    stable=True keeps input order inside every destination (the reference's permuted order); stable=False is the
    faster tile-sorting scatter whose order inside a destination is unspecified (perm still gives the origin).
    Returns (permuted keys, perm or None, counts[p] as a python list)."""
    dev = keys.device.index
    W = 1 if keys.dim() == 1 else keys.shape[1]
    a, n = _dev_keys(keys, W, dev)
    out = torch.empty_like(a)
    perm = torch.empty(n, dtype=torch.int32, device=a.device) if want_perm else None
    counts = np.zeros(p, dtype=np.uint64)
    fn = lib.kh_partition_dev if stable else lib.kh_partition_fast_dev
    check(fn(dev, HASH_KINDS[hash_kind], seed, W, p, a.data_ptr(), n, out.data_ptr(),
             perm.data_ptr() if want_perm else None, counts.ctypes.data), "partition")
    return out, perm, [int(c) for c in counts]


def partition(keys, p, hash_kind="MURMUR64avx", seed=DIST_SEED, device=0):
    """Host-array convenience wrapper of partition_dev: (permuted keys, perm, counts) as numpy."""
    a, W, n = _shape_keys(keys)
    t = torch.from_numpy(a.view(np.int64)).to("cuda:%d" % device)
    out, perm, counts = partition_dev(t, p, hash_kind, seed, True)
    return out.cpu().numpy().view(np.uint64), perm.cpu().numpy().view(np.uint32), counts


def unpermute_dev(vals, perm):
    """out[perm[i]] = vals[i] for uint8 / 32-bit CUDA tensors (query results back to input order)."""
    dev = vals.device.index
    out = torch.empty_like(vals)
    torch.cuda.current_stream(dev).synchronize()
    check(lib.kh_unpermute_dev(dev, vals.element_size(), vals.data_ptr(), perm.data_ptr(), vals.shape[0], out.data_ptr()),
          "unpermute")
    return out


class MultiGpuMap:
    """Single-process map over n GPUs (khd_*): dsc::counting_batched_robinhood_map stand-in.
    Results of count/find come back in input order."""

    def __init__(self, n_gpus, key_words=1, store_hash="MURMUR64avx", dist_hash="MURMUR64avx", reducer="COUNT", canonical_k=0):
        self.W = key_words
        self._h = lib.khd_create(n_gpus, key_words, HASH_KINDS[store_hash], HASH_KINDS[dist_hash], REDUCERS[reducer], canonical_k)
        if not self._h:
            check(-3 if lib.kh_device_count() <= 0 else -1, "khd_create")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib.khd_destroy(h)

    def n_gpus(self):
        return lib.khd_n_gpus(self._h)

    def uses_peer_stores(self):
        return bool(lib.khd_uses_peer_stores(self._h))

    def insert(self, keys, default=1, vals=None):
        a, n = _host_keys(keys, self.W)
        if vals is not None:              # (key, value) pairs: FIRST / LAST maps
            v = np.ascontiguousarray(vals, dtype=np.uint32)
            assert v.shape[0] == n
            return check(lib.khd_insert_pairs_host(self._h, a.ctypes.data, v.ctypes.data, n), "khd_insert_pairs")
        return check(lib.khd_insert_host(self._h, a.ctypes.data, n, default), "khd_insert")

    def set_load_factors(self, min_load, max_load):
        check(lib.khd_set_load_factors(self._h, min_load, max_load), "khd_set_load_factors")

    def count(self, keys):
        a, n = _host_keys(keys, self.W)
        out = np.empty(n, dtype=np.uint8)
        check(lib.khd_count_host(self._h, a.ctypes.data, n, out.ctypes.data), "khd_count")
        return out

    def find(self, keys, nonexistent=0):
        a, n = _host_keys(keys, self.W)
        out = np.empty(n, dtype=np.uint32)
        check(lib.khd_find_host(self._h, a.ctypes.data, n, out.ctypes.data, nonexistent), "khd_find")
        return out

    def erase(self, keys):
        a, n = _host_keys(keys, self.W)
        return check(lib.khd_erase_host(self._h, a.ctypes.data, n), "khd_erase")

    def size(self):
        return lib.khd_size(self._h)

    def local_size(self, rank):
        return lib.khd_local_size(self._h, rank)

    def to_vector_rank(self, rank):
        n = self.local_size(rank)
        ok = np.empty((max(n, 1), self.W), dtype=np.uint64)
        ov = np.empty(max(n, 1), dtype=np.uint32)
        c = min(check(lib.khd_export_rank_host(self._h, rank, ok.ctypes.data, ov.ctypes.data, n), "khd_export"), n)
        return (ok[:c] if self.W > 1 else ok[:c, 0].copy()), ov[:c]


def probe_random_access(buf_bytes, n_access, bytes_per_access=32, mode="read", iters=3, device=0):
    """milliseconds of the random-access roofline probe."""
    ms = lib.kh_probe_random_access(device, buf_bytes, n_access, bytes_per_access, {"read": 0, "rmw": 1, "read_ilp4": 2, "rmw_ilp4": 3}[mode], iters)
    if ms < 0:
        check(-3, "probe_random_access")
    return ms


def probe_stream_copy(nbytes, iters=3, device=0):
    ms = lib.kh_probe_stream_copy(device, nbytes, iters)
    if ms < 0:
        check(-3, "probe_stream_copy")
    return ms
