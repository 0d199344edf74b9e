"""ctypes binding of libkmerhash_b200.so (the C ABI in include/kmerhash_b200.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``make -C kmerhash_b200/csrc``.
There is no CPU fallback: if the shared library is missing the import fails loudly, and every
compute call fails with KH_ERR_CUDA when no CUDA device is visible.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkmerhash_b200.so")

HASH_KINDS = {
    "IDEN": 0, "MURMUR64": 1, "MURMUR64avx": 2, "MURMUR32": 3, "MURMUR32avx": 4,
    "MURMUR32sse": 5, "FARM": 6, "FARM32": 7, "CLHASH": 8, "CRC32C": 9,
    "MURMUR32FINALIZER": 10, "MURMUR_X86": 11,
}
REDUCERS = {"COUNT": 0, "FIRST": 1, "LAST": 2}
MAP_KINDS = {"ROBINHOOD_OFFSETS": 0, "ROBINHOOD": 1, "LINEARPROBE": 2}
DEFAULT_SEED, DEFAULT_CRC_SEED, DIST_SEED = 43, 37, 9876543

KH_OK, KH_ERR_INVALID, KH_ERR_NOMEM, KH_ERR_CUDA, KH_ERR_FULL = 0, -1, -2, -3, -4


class KmerHashError(RuntimeError):
    pass


def build_library(verbose=False):
    import subprocess
    r = subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-j8"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-3000:], r.stderr[-3000:])
    if r.returncode != 0:
        raise KmerHashError("building libkmerhash_b200.so failed")


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "kmerhash_b200: %s is missing. Build it with __graft_entry__.build() "
            "(nvcc, sm_100a). There is no CPU fallback." % LIB_PATH)
    return C.CDLL(LIB_PATH)


lib = _load()

_vp, _u64, _u32, _i64, _int, _dbl = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int64, C.c_int, C.c_double

# name -> (restype, argtypes).  Pointers are void* so numpy (host) and torch (device) addresses both fit.
SIGNATURES = {
    "kh_last_error": (C.c_char_p, []),
    "kh_device_count": (_int, []),
    "kh_launch_count": (_u64, []),
    "kh_reset_launch_count": (None, []),
    "kh_host_alloc": (_vp, [C.c_size_t]),
    "kh_host_free": (None, [_vp]),
    "kh_create": (_vp, [_int, _int, _int, _int, _u32, _int]),
    "kh_destroy": (None, [_vp]),
    "kh_set_load_factors": (_int, [_vp, _dbl, _dbl]),
    "kh_set_lookahead": (_int, [_vp, _int, _int]),
    "kh_set_geometry": (_int, [_vp, _int, _int]),
    "kh_stream": (_vp, [_vp]),
    "kh_kernel_timing": (_int, [_vp, _int]),
    "kh_kernel_time": (_int, [_vp, _int, _vp, _vp]),
    "kh_reserve": (_int, [_vp, _u64]),
    "kh_rehash": (_int, [_vp, _u64]),
    "kh_capacity": (_u64, [_vp]),
    "kh_size": (_u64, [_vp]),
    "kh_load_factor": (_dbl, [_vp]),
    "kh_clear": (_int, [_vp]),
    "kh_device": (_int, [_vp]),
    "kh_key_words": (_int, [_vp]),
    "kh_estimate_distinct_host": (_dbl, [_vp, _vp, _u64, _vp]),
    "kh_estimate_distinct_dev": (_dbl, [_vp, _vp, _u64, _vp]),
    "kh_insert_host": (_i64, [_vp, _vp, _vp, _u64, _u32, _int]),
    "kh_insert_dev": (_i64, [_vp, _vp, _vp, _u64, _u32, _int]),
    "kh_set_insert_window": (_int, [_vp, _int, _u64, _u64]),
    "kh_insert_pairs_host": (_i64, [_vp, _vp, _u64, _u32, _u32, _u32, _int]),
    "kh_count_host": (_i64, [_vp, _vp, _u64, _vp]),
    "kh_count_dev": (_i64, [_vp, _vp, _u64, _vp]),
    "kh_find_host": (_i64, [_vp, _vp, _u64, _vp, _u32]),
    "kh_find_dev": (_i64, [_vp, _vp, _u64, _vp, _u32]),
    "kh_find_existing_host": (_i64, [_vp, _vp, _u64, _vp, _vp]),
    "kh_find_existing_dev": (_i64, [_vp, _vp, _u64, _vp, _vp]),
    "kh_update_host": (_i64, [_vp, _vp, _vp, _u64]),
    "kh_update_dev": (_i64, [_vp, _vp, _vp, _u64]),
    "kh_erase_host": (_i64, [_vp, _vp, _u64]),
    "kh_erase_dev": (_i64, [_vp, _vp, _u64]),
    "kh_export_host": (_i64, [_vp, _vp, _vp, _u64]),
    "kh_export_dev": (_i64, [_vp, _vp, _vp, _u64]),
    "kh_hash_host": (_int, [_int, _int, _u32, _int, _vp, _u64, _vp]),
    "kh_hash_dev": (_int, [_int, _int, _u32, _int, _vp, _u64, _vp]),
    "kh_kmers_from_reads_host": (_i64, [_int, _vp, _vp, _u64, _int, _int, _vp, _u64]),
    "kh_kmer_words_alpha": (_int, [_int, _int]),
    "kh_kmers_from_reads_alpha_dev": (_i64, [_int, _int, _vp, _vp, _u64, _int, _int, _vp, _u64]),
    "kh_kmers_from_reads_alpha_host": (_i64, [_int, _int, _vp, _vp, _u64, _int, _int, _vp, _u64]),
    "kh_canonicalize_alpha_host": (_int, [_int, _int, _int, _vp, _u64, _vp]),
    "kh_revcomp_alpha_host": (_int, [_int, _int, _int, _vp, _u64, _vp]),
    "kh_kmers_from_reads_dev": (_i64, [_int, _vp, _vp, _u64, _int, _int, _vp, _u64]),
    "kh_canonicalize_dev": (_int, [_int, _int, _vp, _u64, _vp]),
    "kh_canonicalize_host": (_int, [_int, _int, _vp, _u64, _vp]),
    "kh_revcomp_host": (_int, [_int, _int, _vp, _u64, _vp]),
    "kh_partition_dev": (_int, [_int, _int, _u32, _int, _u32, _vp, _u64, _vp, _vp, _vp]),
    "khx_meta_words": (_int, []),
    "khx_peer_create": (_vp, [_int, _u32, _u32, _u64, _int]),
    "khx_peer_destroy": (None, [_vp]),
    "khx_peer_capacity": (_u64, [_vp]),
    "khx_peer_recv_ptr": (_vp, [_vp]),
    "khx_peer_handle": (_int, [_vp, _vp]),
    "khx_peer_open": (_int, [_vp, _u32, _vp]),
    "khx_peer_set": (_int, [_vp, _u32, _vp]),
    "khx_plan_dev": (_int, [_vp, _int, _u32, _u32, _vp, _u64, _vp]),
    "khx_scatter_peers_dev": (_int, [_vp, _int, _u32, _vp, _u64, _vp, _vp, _vp]),
    "khx_plan_scatter_slabs_dev": (_int, [_vp, _int, _u32, _vp, _u64, _vp, _u64, _int, _vp]),
    "khx_insert_recv_slabs_dev": (_i64, [_vp, _vp, _u64, _int, _vp, _vp, _u32]),
    "khx_push_meta_sync_dev": (_int, [_vp, _vp, _int, _vp, _u64, _vp]),
    "khx_peer_meta_ptr": (_vp, [_vp, _int]),
    "khx_scatter_slabs_dev": (_int, [_vp, _int, _u32, _vp, _u64, _vp, _u64, _int, _vp, _int]),
    "khx_insert_pipelined_dev": (_i64, [_vp, _vp, _int, _u32, _vp, _u64, _u64, _int, _int, _u64, _u32, _vp, _vp]),
    "khx_insert_ce_dev": (_i64, [_vp, _vp, _int, _u32, _vp, _u64, _u64, _int, _int, _u64, _u32, _vp, _vp]),
    "khx_kmers_scatter_slabs_dev": (_int, [_vp, _vp, _vp, _u64, _int, _int, _int, _u32, _vp, _u64, _int, _vp]),
    "khx_peer_close_imports": (None, [_vp]),
    "khx_peer_poisoned": (_int, [_vp]),
    "khx_peer_sync_dev": (_int, [_vp, _vp, _u64, _vp, _vp]),
    "khx_peer_create_q": (_vp, [_int, _u32, _u32, _u64, _int]),
    "khx_query_scatter_slabs_dev": (_int, [_vp, _int, _u32, _vp, _u64, _vp, _u64, _vp, _vp]),
    "khx_query_slabs_dev": (_i64, [_vp, _vp, _int, _u64, _vp, _u32]),
    "khx_unpermute_slabs_dev": (_int, [_vp, _vp, _int, _u64, _vp, _vp, _u64, _vp]),
    "khx_plan_scatter_dev": (_int, [_vp, _int, _u32, _u32, _vp, _u64, _vp, _vp, _vp]),
    "khx_insert_recv_dev": (_i64, [_vp, _vp, _u64, _vp, _u32, _u32]),
    "kh_partition_fast_dev": (_int, [_int, _int, _u32, _int, _u32, _vp, _u64, _vp, _vp, _vp]),
    "kh_partition_fast_noalloc_dev": (_int, [_int, _int, _u32, _int, _u32, _vp, _u64, _vp, _vp, _vp]),
    "kh_rank_host": (_int, [_int, _int, _u32, _int, _u32, _vp, _u64, _vp]),
    "kh_unpermute_dev": (_int, [_int, _int, _vp, _vp, _u64, _vp]),
    "khd_create": (_vp, [_int, _int, _int, _int, _int, _int]),
    "khd_destroy": (None, [_vp]),
    "khd_n_gpus": (_int, [_vp]),
    "khd_uses_peer_stores": (_int, [_vp]),
    "khd_insert_host": (_i64, [_vp, _vp, _u64, _u32]),
    "khd_insert_pairs_host": (_i64, [_vp, _vp, _vp, _u64]),
    "khd_set_load_factors": (_int, [_vp, _dbl, _dbl]),
    "khd_count_host": (_i64, [_vp, _vp, _u64, _vp]),
    "khd_find_host": (_i64, [_vp, _vp, _u64, _vp, _u32]),
    "khd_erase_host": (_i64, [_vp, _vp, _u64]),
    "khd_size": (_u64, [_vp]),
    "khd_local_size": (_u64, [_vp, _int]),
    "khd_export_rank_host": (_i64, [_vp, _int, _vp, _vp, _u64]),
    "kh_probe_random_access": (_dbl, [_int, _u64, _u64, _int, _int, _int]),
    "kh_probe_stream_copy": (_dbl, [_int, _u64, _int]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)          # AttributeError here = the .so does not export what the header declares
    _fn.restype = _res
    _fn.argtypes = _args


def last_error():
    msg = lib.kh_last_error()
    return msg.decode() if msg else ""


def check(rc, what):
    """Negative status -> exception of the type the reference would throw."""
    if rc is None or rc >= 0:
        return rc
    msg = "%s failed (%d): %s" % (what, rc, last_error())
    if rc == KH_ERR_INVALID:
        raise ValueError(msg)          # std::invalid_argument
    if rc == KH_ERR_NOMEM:
        raise MemoryError(msg)         # std::length_error (KH/mem_utils.hpp:47)
    raise KmerHashError(msg)           # std::logic_error
