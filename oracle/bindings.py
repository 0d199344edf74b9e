"""ctypes bindings of the two checkers (see oracle/__init__.py).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# same numbering as include/kmerhash_b200.h (kh_hash_kind)
HASH_KINDS = {
    "IDEN": 0, "MURMUR64": 1, "MURMUR64avx": 2, "MURMUR32": 3, "MURMUR32avx": 4,
    "MURMUR32sse": 5, "FARM": 6, "FARM32": 7, "CLHASH": 8, "CRC32C": 9,
    "MURMUR32FINALIZER": 10, "MURMUR_X86": 11,
}
MAP_KINDS = {"OFFSETS_COUNT": 0, "OFFSETS_FIRST": 1, "OFFSETS_LAST": 2, "ROBINHOOD": 3, "LINEARPROBE": 4}

_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")


def build(verbose=False):
    """(Re)build the checkers.  _ref is only rebuilt where /root/reference exists."""
    r = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("oracle build failed")


def _kw(keys):
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    if keys.ndim == 1:
        return keys, 1, keys.shape[0]
    return keys, keys.shape[1], keys.shape[0]


class _Port:
    """Plain-C restatement (kmerhash_oracle.c)."""

    def __init__(self):
        path = os.path.join(HERE, "libkmerhash_oracle.so")
        if not os.path.exists(path):
            build()
        L = self.lib = C.CDLL(path)
        L.orc_hash.argtypes = [C.c_int, C.c_uint64, C.c_int, _u64p, C.c_size_t, _u64p]
        L.orc_hash.restype = C.c_int
        L.orc_rand_words.argtypes = [C.c_uint32, C.c_int, _u64p]
        L.orc_kmer_windows.argtypes = [C.c_int, C.c_char_p, C.c_size_t, _u64p]
        L.orc_kmer_windows.restype = C.c_long
        L.orc_revcomp.argtypes = [C.c_int, _u64p, C.c_size_t, _u64p]
        L.orc_canonical.argtypes = [C.c_int, _u64p, C.c_size_t, _u64p]
        L.orc_kmers_from_reads.argtypes = [C.c_char_p, _u64p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        L.orc_kmers_from_reads.restype = C.c_long
        L.orc_rank.argtypes = [C.c_int, C.c_int, _u64p, C.c_size_t, C.c_uint32, _u32p]
        L.orc_rank.restype = C.c_int
        L.orc_kmer_words_alpha.argtypes = [C.c_int, C.c_int]
        L.orc_kmer_windows_alpha.argtypes = [C.c_int, C.c_int, C.c_char_p, C.c_size_t, _u64p]
        L.orc_kmer_windows_alpha.restype = C.c_long
        L.orc_revcomp_alpha.argtypes = [C.c_int, C.c_int, _u64p, C.c_size_t, _u64p]
        L.orc_canonical_alpha.argtypes = [C.c_int, C.c_int, _u64p, C.c_size_t, _u64p]
        L.orc_kmers_from_reads_alpha.argtypes = [C.c_int, C.c_char_p, _u64p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        L.orc_kmers_from_reads_alpha.restype = C.c_long
        L.orc_hll_update.argtypes = [_u8p, _u64p, C.c_size_t, C.c_int]
        L.orc_hll_estimate.argtypes = [_u8p, C.c_int]
        L.orc_hll_estimate.restype = C.c_double
        L.orc_map_create.argtypes = [C.c_int, C.c_int]
        L.orc_map_create.restype = C.c_void_p
        L.orc_map_destroy.argtypes = [C.c_void_p]
        L.orc_map_insert.argtypes = [C.c_void_p, _u64p, C.c_void_p, C.c_uint32, C.c_size_t]
        L.orc_map_insert.restype = C.c_size_t
        L.orc_map_count.argtypes = [C.c_void_p, _u64p, C.c_size_t, _u8p]
        L.orc_map_count.restype = C.c_size_t
        L.orc_map_find.argtypes = [C.c_void_p, _u64p, C.c_size_t, _u32p, C.c_uint32]
        L.orc_map_find.restype = C.c_size_t
        L.orc_map_find_existing.argtypes = [C.c_void_p, _u64p, C.c_size_t, _u64p, _u32p]
        L.orc_map_find_existing.restype = C.c_size_t
        L.orc_map_erase.argtypes = [C.c_void_p, _u64p, C.c_size_t]
        L.orc_map_erase.restype = C.c_size_t
        L.orc_map_size.argtypes = [C.c_void_p]
        L.orc_map_size.restype = C.c_size_t
        L.orc_map_to_vector.argtypes = [C.c_void_p, _u64p, _u32p, C.c_size_t]
        L.orc_map_to_vector.restype = C.c_size_t
        L.orc_generate_input.argtypes = [C.c_int, C.c_size_t, C.c_size_t, C.c_int, _u64p, _u32p]
        L.orc_generate_input.restype = C.c_size_t
        L.orc_next_pow2.argtypes = [C.c_uint64]
        L.orc_next_pow2.restype = C.c_uint64

    def hash(self, kind, keys, seed=43):
        keys, w, n = _kw(keys)
        out = np.empty(n, dtype=np.uint64)
        if self.lib.orc_hash(HASH_KINDS[kind], seed, w, keys, n, out) != 0:
            raise ValueError("unsupported hash")
        return out

    def rand_words(self, seed, n):
        out = np.empty(n, dtype=np.uint64)
        self.lib.orc_rand_words(seed, n, out)
        return out

    @staticmethod
    def words(k):
        return (2 * k + 63) // 64

    def kmer_windows(self, k, s):
        if isinstance(s, str):
            s = s.encode()
        w = self.words(k)
        out = np.zeros((max(len(s), 1), w), dtype=np.uint64)
        n = self.lib.orc_kmer_windows(k, s, len(s), out)
        return out[:n].reshape(n, w) if w > 1 else out[:n, 0].copy()

    def revcomp(self, k, kmers):
        kmers = np.ascontiguousarray(kmers, dtype=np.uint64)
        out = np.empty_like(kmers)
        self.lib.orc_revcomp(k, kmers, kmers.shape[0], out)
        return out

    def canonical(self, k, kmers):
        kmers = np.ascontiguousarray(kmers, dtype=np.uint64)
        out = np.empty_like(kmers)
        self.lib.orc_canonical(k, kmers, kmers.shape[0], out)
        return out

    def kmers_from_reads(self, bases, offsets, k, canonical):
        if isinstance(bases, np.ndarray):
            bases = bases.tobytes()
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n_reads = offsets.shape[0] - 1
        n = self.lib.orc_kmers_from_reads(bases, offsets, n_reads, k, int(canonical), None)
        w = self.words(k)
        out = np.zeros((max(n, 1), w), dtype=np.uint64)
        self.lib.orc_kmers_from_reads(bases, offsets, n_reads, k, int(canonical), out.ctypes.data_as(C.c_void_p))
        return out[:n] if w > 1 else out[:n, 0].copy()

    # ---- other alphabets: alpha = 5 (DNA5, 3 bits) / 16 (DNA16, 4 bits) / 4 (DNA) ----
    def words_alpha(self, alpha, k):
        return self.lib.orc_kmer_words_alpha(alpha, k)

    def kmer_windows_alpha(self, alpha, k, s):
        if isinstance(s, str):
            s = s.encode()
        w = self.words_alpha(alpha, k)
        out = np.zeros((max(len(s), 1), w), dtype=np.uint64)
        n = self.lib.orc_kmer_windows_alpha(alpha, k, s, len(s), out)
        return out[:n].reshape(n, w) if w > 1 else out[:n, 0].copy()

    def revcomp_alpha(self, alpha, k, kmers):
        kmers = np.ascontiguousarray(kmers, dtype=np.uint64)
        out = np.empty_like(kmers)
        assert self.lib.orc_revcomp_alpha(alpha, k, kmers, kmers.shape[0], out) == 0
        return out

    def canonical_alpha(self, alpha, k, kmers):
        kmers = np.ascontiguousarray(kmers, dtype=np.uint64)
        out = np.empty_like(kmers)
        assert self.lib.orc_canonical_alpha(alpha, k, kmers, kmers.shape[0], out) == 0
        return out

    def kmers_from_reads_alpha(self, alpha, bases, offsets, k, canonical):
        if isinstance(bases, np.ndarray):
            bases = bases.tobytes()
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n_reads = offsets.shape[0] - 1
        n = self.lib.orc_kmers_from_reads_alpha(alpha, bases, offsets, n_reads, k, int(canonical), None)
        w = self.words_alpha(alpha, k)
        out = np.zeros((max(n, 1), w), dtype=np.uint64)
        self.lib.orc_kmers_from_reads_alpha(alpha, bases, offsets, n_reads, k, int(canonical), out.ctypes.data_as(C.c_void_p))
        return out[:n] if w > 1 else out[:n, 0].copy()

    def rank(self, kind, keys, p):
        keys, w, n = _kw(keys)
        out = np.empty(n, dtype=np.uint32)
        if self.lib.orc_rank(HASH_KINDS[kind], w, keys, n, p, out) != 0:
            raise ValueError("bad rank args")
        return out

    def hll_estimate(self, hashes, hash_bits=64):
        regs = np.zeros(4096, dtype=np.uint8)
        hashes = np.ascontiguousarray(hashes, dtype=np.uint64)
        self.lib.orc_hll_update(regs, hashes, hashes.shape[0], hash_bits)
        return self.lib.orc_hll_estimate(regs, hash_bits), regs

    def generate_input(self, k, count, repeats=10, canonical=False):
        w = self.words(k)
        keys = np.empty((count, w), dtype=np.uint64)
        vals = np.empty(count, dtype=np.uint32)
        n = self.lib.orc_generate_input(k, count, repeats, int(canonical), keys, vals)
        keys = keys[:n]
        return (keys if w > 1 else keys[:, 0].copy()), vals[:n]

    def Map(self, key_words=1, reducer="COUNT"):
        return _PortMap(self.lib, key_words, {"COUNT": 0, "FIRST": 1, "LAST": 2}[reducer])


class _PortMap:
    def __init__(self, lib, key_words, reducer):
        self.lib, self.w = lib, key_words
        self.h = lib.orc_map_create(key_words, reducer)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.orc_map_destroy(self.h)
            self.h = None

    def _k(self, keys):
        keys, w, n = _kw(keys)
        assert w == self.w
        return keys, n

    def insert(self, keys, vals=None, default=1):
        keys, n = self._k(keys)
        vp = None
        if vals is not None:
            vals = np.ascontiguousarray(vals, dtype=np.uint32)
            vp = vals.ctypes.data_as(C.c_void_p)
        return self.lib.orc_map_insert(self.h, keys, vp, default, n)

    def count(self, keys):
        keys, n = self._k(keys)
        out = np.empty(n, dtype=np.uint8)
        self.lib.orc_map_count(self.h, keys, n, out)
        return out

    def find(self, keys, nonexistent=0):
        keys, n = self._k(keys)
        out = np.empty(n, dtype=np.uint32)
        self.lib.orc_map_find(self.h, keys, n, out, nonexistent)
        return out

    def find_existing(self, keys):
        keys, n = self._k(keys)
        ok = np.empty((max(n, 1), self.w), dtype=np.uint64)
        ov = np.empty(max(n, 1), dtype=np.uint32)
        c = self.lib.orc_map_find_existing(self.h, keys, n, ok, ov)
        return (ok[:c] if self.w > 1 else ok[:c, 0].copy()), ov[:c]

    def erase(self, keys):
        keys, n = self._k(keys)
        return self.lib.orc_map_erase(self.h, keys, n)

    def size(self):
        return self.lib.orc_map_size(self.h)

    def to_vector(self):
        n = self.size()
        ok = np.empty((max(n, 1), self.w), dtype=np.uint64)
        ov = np.empty(max(n, 1), dtype=np.uint32)
        self.lib.orc_map_to_vector(self.h, ok, ov, n)
        return (ok[:n] if self.w > 1 else ok[:n, 0].copy()), ov[:n]


class _Ref:
    """The reference's own implementation (oracle/_ref/libkmerhash_ref.so)."""

    def __init__(self, path):
        L = self.lib = C.CDLL(path)
        L.ref_hash.argtypes = [C.c_int, C.c_uint64, C.c_int, _u64p, C.c_size_t, _u64p, C.c_int]
        L.ref_hash.restype = C.c_int
        L.ref_kmer_windows.argtypes = [C.c_int, C.c_char_p, C.c_size_t, _u64p]
        L.ref_kmer_windows.restype = C.c_long
        L.ref_revcomp.argtypes = [C.c_int, _u64p, C.c_size_t, _u64p]
        L.ref_revcomp.restype = C.c_int
        L.ref_canonical.argtypes = [C.c_int, _u64p, C.c_size_t, _u64p]
        L.ref_canonical.restype = C.c_int
        L.ref_kmer_words.argtypes = [C.c_int]
        if hasattr(L, "ref_kmer_windows_alpha"):
            L.ref_kmer_windows_alpha.argtypes = [C.c_int, C.c_int, C.c_char_p, C.c_size_t, _u64p]
            L.ref_kmer_windows_alpha.restype = C.c_long
            L.ref_revcomp_alpha.argtypes = [C.c_int, C.c_int, _u64p, C.c_size_t, _u64p]
            L.ref_canonical_alpha.argtypes = [C.c_int, C.c_int, _u64p, C.c_size_t, _u64p]
            L.ref_kmer_words_alpha.argtypes = [C.c_int, C.c_int]
        L.ref_kmer_words.restype = C.c_int
        L.ref_map_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]
        L.ref_map_create.restype = C.c_void_p
        L.ref_map_destroy.argtypes = [C.c_void_p]
        L.ref_map_insert_keys.argtypes = [C.c_void_p, _u64p, C.c_size_t, C.c_uint32, C.c_int]
        L.ref_map_insert_pairs.argtypes = [C.c_void_p, _u64p, _u32p, C.c_size_t, C.c_int]
        L.ref_map_count.argtypes = [C.c_void_p, _u64p, C.c_size_t, _u8p]
        L.ref_map_count.restype = C.c_size_t
        L.ref_map_find.argtypes = [C.c_void_p, _u64p, C.c_size_t, _u32p, C.c_uint32]
        L.ref_map_find.restype = C.c_size_t
        L.ref_map_find_existing.argtypes = [C.c_void_p, _u64p, C.c_size_t, _u64p, _u32p]
        L.ref_map_find_existing.restype = C.c_size_t
        L.ref_map_erase.argtypes = [C.c_void_p, _u64p, C.c_size_t]
        L.ref_map_erase.restype = C.c_size_t
        L.ref_map_to_vector.argtypes = [C.c_void_p, _u64p, _u32p, C.c_size_t]
        L.ref_map_to_vector.restype = C.c_size_t
        L.ref_map_size.argtypes = [C.c_void_p]
        L.ref_map_size.restype = C.c_size_t
        L.ref_map_capacity.argtypes = [C.c_void_p]
        L.ref_map_capacity.restype = C.c_size_t
        L.ref_map_reserve.argtypes = [C.c_void_p, C.c_size_t]
        L.ref_map_hll_estimate.argtypes = [C.c_void_p, _u64p, C.c_size_t]
        L.ref_map_hll_estimate.restype = C.c_double
        L.ref_rank.argtypes = [C.c_int, C.c_int, _u64p, C.c_size_t, C.c_uint32, _u32p]
        L.ref_rank.restype = C.c_int
        L.ref_generate_input.argtypes = [C.c_int, C.c_size_t, C.c_size_t, C.c_int, _u64p, _u32p]
        L.ref_generate_input.restype = C.c_size_t
        L.ref_rand_words.argtypes = [C.c_uint, C.c_int, _u64p]
        L.ref_baseline_run.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, _u64p, C.c_size_t,
                                       _u64p, C.c_size_t, C.c_int, _f64p, _u64p]
        L.ref_baseline_run.restype = C.c_int

    def hash(self, kind, keys, seed=43, batch=False):
        keys, w, n = _kw(keys)
        out = np.empty(n, dtype=np.uint64)
        if self.lib.ref_hash(HASH_KINDS[kind], seed, w, keys, n, out, int(batch)) != 0:
            raise ValueError("unsupported hash")
        return out

    def rand_words(self, seed, n):
        out = np.empty(n, dtype=np.uint64)
        self.lib.ref_rand_words(seed, n, out)
        return out

    def kmer_windows(self, k, s):
        if isinstance(s, str):
            s = s.encode()
        w = self.lib.ref_kmer_words(k)
        out = np.zeros((max(len(s), 1), w), dtype=np.uint64)
        n = self.lib.ref_kmer_windows(k, s, len(s), out)
        return out[:n] if w > 1 else out[:n, 0].copy()

    def revcomp(self, k, kmers):
        kmers = np.ascontiguousarray(kmers, dtype=np.uint64)
        out = np.empty_like(kmers)
        assert self.lib.ref_revcomp(k, kmers, kmers.shape[0], out) == 0
        return out

    def canonical(self, k, kmers):
        kmers = np.ascontiguousarray(kmers, dtype=np.uint64)
        out = np.empty_like(kmers)
        assert self.lib.ref_canonical(k, kmers, kmers.shape[0], out) == 0
        return out

    # ---- other alphabets (compiled reference: Kmer<k, DNA5 / DNA16, uint64_t> for the (alpha, k) pairs instantiated) ----
    def kmer_windows_alpha(self, alpha, k, s):
        if isinstance(s, str):
            s = s.encode()
        w = self.lib.ref_kmer_words_alpha(alpha, k)
        if w < 0:
            raise ValueError("(alphabet, k) not instantiated in oracle/_ref")
        out = np.zeros((max(len(s), 1), w), dtype=np.uint64)
        n = self.lib.ref_kmer_windows_alpha(alpha, k, s, len(s), out)
        return out[:n] if w > 1 else out[:n, 0].copy()

    def revcomp_alpha(self, alpha, k, kmers):
        kmers = np.ascontiguousarray(kmers, dtype=np.uint64)
        out = np.empty_like(kmers)
        assert self.lib.ref_revcomp_alpha(alpha, k, kmers, kmers.shape[0], out) == 0
        return out

    def canonical_alpha(self, alpha, k, kmers):
        kmers = np.ascontiguousarray(kmers, dtype=np.uint64)
        out = np.empty_like(kmers)
        assert self.lib.ref_canonical_alpha(alpha, k, kmers, kmers.shape[0], out) == 0
        return out

    def rank(self, kind, keys, p):
        keys, w, n = _kw(keys)
        out = np.empty(n, dtype=np.uint32)
        if self.lib.ref_rank(HASH_KINDS[kind], w, keys, n, p, out) != 0:
            raise ValueError("bad rank args")
        return out

    def generate_input(self, key_words, count, repeats=10, canonical=False):
        keys = np.empty((count, key_words), dtype=np.uint64)
        vals = np.empty(count, dtype=np.uint32)
        n = self.lib.ref_generate_input(key_words, count, repeats, int(canonical), keys, vals)
        keys = keys[:n]
        return (keys if key_words > 1 else keys[:, 0].copy()), vals[:n]

    def Map(self, map_kind="OFFSETS_COUNT", key_words=1, hash_kind="MURMUR64avx", min_load=0.35, max_load=0.8):
        return _RefMap(self.lib, MAP_KINDS[map_kind], key_words, HASH_KINDS[hash_kind], min_load, max_load)

    def baseline_run(self, keys, queries, map_kind="OFFSETS_COUNT", hash_kind="MURMUR64avx", threads=1,
                     min_load=0.35, max_load=0.8):
        keys, w, n = _kw(keys)
        queries, wq, nq = _kw(queries)
        secs = np.zeros(7, dtype=np.float64)
        cnts = np.zeros(5, dtype=np.uint64)
        rc = self.lib.ref_baseline_run(MAP_KINDS[map_kind], w, HASH_KINDS[hash_kind], min_load, max_load,
                                       keys, n, queries, nq, threads, secs, cnts)
        if rc != 0:
            raise RuntimeError("ref_baseline_run failed")
        names = ["partition", "estimate", "insert", "find", "count", "erase", "count2"]
        return dict(zip(names, secs.tolist())), dict(zip(["distinct", "found", "counted", "erased", "counted2"], cnts.tolist()))


class _RefMap:
    def __init__(self, lib, mk, w, hk, minl, maxl):
        self.lib, self.w = lib, w
        self.h = lib.ref_map_create(mk, w, hk, minl, maxl)
        if not self.h:
            raise ValueError("map/hash combination not instantiated in oracle/_ref")

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.ref_map_destroy(self.h)
            self.h = None

    def _k(self, keys):
        keys, w, n = _kw(keys)
        assert w == self.w
        return keys, n

    def insert(self, keys, vals=None, default=1, estimate=True):
        keys, n = self._k(keys)
        if vals is None:
            self.lib.ref_map_insert_keys(self.h, keys, n, default, int(estimate))
        else:
            self.lib.ref_map_insert_pairs(self.h, keys, np.ascontiguousarray(vals, dtype=np.uint32), n, int(estimate))

    def count(self, keys):
        keys, n = self._k(keys)
        out = np.zeros(n, dtype=np.uint8)
        self.lib.ref_map_count(self.h, keys, n, out)
        return out

    def find(self, keys, nonexistent=0):
        keys, n = self._k(keys)
        out = np.zeros(n, dtype=np.uint32)
        self.lib.ref_map_find(self.h, keys, n, out, nonexistent)
        return out

    def find_existing(self, keys):
        keys, n = self._k(keys)
        ok = np.empty((max(n, 1), self.w), dtype=np.uint64)
        ov = np.empty(max(n, 1), dtype=np.uint32)
        c = self.lib.ref_map_find_existing(self.h, keys, n, ok, ov)
        return (ok[:c] if self.w > 1 else ok[:c, 0].copy()), ov[:c]

    def erase(self, keys):
        keys, n = self._k(keys)
        return self.lib.ref_map_erase(self.h, keys, n)

    def size(self):
        return self.lib.ref_map_size(self.h)

    def capacity(self):
        return self.lib.ref_map_capacity(self.h)

    def reserve(self, n):
        self.lib.ref_map_reserve(self.h, n)

    def hll_estimate(self, keys):
        keys, n = self._k(keys)
        return self.lib.ref_map_hll_estimate(self.h, keys, n)

    def to_vector(self):
        n = self.size()
        ok = np.empty((max(n, 1), self.w), dtype=np.uint64)
        ov = np.empty(max(n, 1), dtype=np.uint32)
        self.lib.ref_map_to_vector(self.h, ok, ov, n)
        return (ok[:n] if self.w > 1 else ok[:n, 0].copy()), ov[:n]


def _load_ref():
    path = os.path.join(HERE, "_ref", "libkmerhash_ref.so")
    if not os.path.exists(path):
        return None
    try:
        return _Ref(path)
    except OSError:
        return None


port = _Port()
ref = _load_ref()
