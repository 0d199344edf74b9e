"""kmerhash_b200 -- B200-native k-mer hash-table engine (hot path of ParBLiSS/kmerhash).

CUDA kernels + C ABI live in csrc/ (built into libkmerhash_b200.so); this package is the thin
host-side mirror of the reference's container interface.  No CPU fallback exists: importing
without the built library raises ImportError, computing without a GPU raises KmerHashError.
"""
from ._lib import (lib, LIB_PATH, HASH_KINDS, REDUCERS, MAP_KINDS, DEFAULT_SEED, DEFAULT_CRC_SEED, DIST_SEED,
                   KmerHashError, build_library, last_error)
from .hashmap import (HashMap, hashmap_robinhood_offsets_count, hashmap_robinhood_offsets,
                      hashmap_robinhood_offsets_reduction, hashmap_robinhood_doubling, hashmap_linearprobe_doubling)
from .ops import (kmer_words, hash_batch, kmers_from_reads, kmers_from_reads_dev, canonicalize, canonicalize_alpha, revcomp, rank_of,
                  partition, partition_dev, unpermute_dev, MultiGpuMap, probe_random_access, probe_stream_copy)


def device_count():
    return lib.kh_device_count()


def launch_count():
    return lib.kh_launch_count()


def reset_launch_count():
    lib.kh_reset_launch_count()
