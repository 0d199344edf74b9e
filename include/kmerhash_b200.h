/* kmerhash_b200.h -- C ABI of the B200-native k-mer hash-table engine.
 *
 * This is the drop-in boundary for the hot path of ParBLiSS/kmerhash:
 * batch insert / find / count / erase of 2-bit packed k-mer keys with count values in an
 * open-addressing table, k-mer generation in front of it and hash-partitioning for the
 * multi-GPU exchange behind it.  The reference has no FFI layer; its "operator API" is the
 * C++ template interface of its containers.  Each entry point below names the reference
 * interface it replaces (paths relative to the reference tree; KH = include/kmerhash,
 * KI = ext/kmerind/src).  include/kmerhash_b200.hpp wraps this ABI in classes with the
 * reference's own names and signatures; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++/torch types.
 *   - keys are arrays of uint64_t words, `key_words` per key (1: k <= 32, 2: k <= 64),
 *     word 0 least significant -- the in-memory layout of bliss::common::Kmer<K,DNA,uint64_t>
 *     (KI/common/kmer.hpp:115-170).
 *   - every `*_host` entry point takes HOST pointers and moves the data itself (chunked,
 *     overlapped with the kernels); every `*_dev` entry point takes DEVICE pointers on the
 *     table's device and is asynchronous only inside the call (it returns after the work
 *     is complete unless noted).
 *   - return value: >= 0 result / 0 ok; < 0 error (kh_last_error() has the text).
 *     KH_ERR_INVALID maps to std::invalid_argument, KH_ERR_NOMEM to std::length_error
 *     (KH/mem_utils.hpp:47), KH_ERR_FULL/KH_ERR_CUDA to std::logic_error
 *     (KH/hashmap_linearprobe.hpp:503) in the C++ facade.
 *   - there is NO CPU fallback: without a CUDA device every call fails with KH_ERR_CUDA.
 */
#ifndef KMERHASH_B200_H_
#define KMERHASH_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KH_OK            0
#define KH_ERR_INVALID  -1
#define KH_ERR_NOMEM    -2
#define KH_ERR_CUDA     -3
#define KH_ERR_FULL     -4

/* Hash functor selector = the reference's -DpStoreHash / -DpDistHash choices
 * (benchmark/BenchmarkHashTables.cpp:131-166; functors in KH/hash_new.hpp:136-351,
 * KH/murmurhash3_64_avx.hpp:1553, KH/murmurhash3_32_avx.hpp:1252, KH/crc32c_sse.hpp:121).
 * Values are bit-exact to the reference functor of the same name. */
typedef enum {
  KH_HASH_IDEN        = 0,  /* fsc::hash::identity: word 0 of the key */
  KH_HASH_MURMUR64    = 1,  /* fsc::hash::murmur: MurmurHash3_x64_128, h[0] */
  KH_HASH_MURMUR64AVX = 2,  /* fsc::hash::murmur3avx64 == murmur_x86: MurmurHash3_x86_128, low 64 bits */
  KH_HASH_MURMUR32    = 3,  /* fsc::hash::murmur32: MurmurHash3_x86_32 */
  KH_HASH_MURMUR32AVX = 4,  /* fsc::hash::murmur3avx32 (== murmur32) */
  KH_HASH_MURMUR32SSE = 5,  /* fsc::hash::murmur3sse32 (== murmur32) */
  KH_HASH_FARM        = 6,  /* fsc::hash::farm: farmhash Hash64WithSeed */
  KH_HASH_FARM32      = 7,  /* fsc::hash::farm32: low 32 bits of farm (SSE4.2+AES build) */
  KH_HASH_CLHASH      = 8,  /* fsc::hash::clhash: key table from glibc srand(seed)/rand() */
  KH_HASH_CRC32C      = 9,  /* fsc::hash::crc32c: crc32q chain started at seed */
  KH_HASH_MURMUR32FIN = 10, /* murmur3finalizer_avx32: keys <= 4 bytes only -> rejected for k-mers */
  KH_HASH_MURMUR_X86  = 11  /* fsc::hash::murmur_x86 (== MURMUR64AVX) */
} kh_hash_kind;

/* Reducer = the reference's value policy on a repeated key
 * (KH/robinhood_offset_hashmap_ptr.hpp:85-97,4094,4099). */
typedef enum {
  KH_REDUCE_COUNT = 0, /* std::plus: v = v_old + v_new  (hashmap_robinhood_offsets_count) */
  KH_REDUCE_FIRST = 1, /* DiscardReducer: first inserted value kept (hashmap_robinhood_offsets, RH, LP maps) */
  KH_REDUCE_LAST  = 2  /* ReplaceReducer: last inserted value kept */
} kh_reducer;

/* Which reference container the table stands in for.  All three share the GPU engine
 * (bucketised open addressing); the kind selects the reference's defaults
 * (initial capacity 128; load factors: offsets 0.2/0.8, robinhood 0.4/0.9, linearprobe
 * 0.2/0.6 -- KH/robinhood_offset_hashmap_ptr.hpp:407-425, KH/hashmap_robinhood.hpp:218-233,
 * KH/hashmap_linearprobe.hpp:191-206) and the default reducer semantics. */
typedef enum {
  KH_MAP_ROBINHOOD_OFFSETS = 0,
  KH_MAP_ROBINHOOD         = 1,
  KH_MAP_LINEARPROBE       = 2
} kh_map_kind;

#define KH_DEFAULT_SEED       43u        /* store-hash seed, KH/hash_new.hpp */
#define KH_DEFAULT_CRC_SEED   37u        /* KH/crc32c_sse.hpp:239 */
#define KH_DIST_SEED          9876543u   /* KH/distributed_batched_robinhood_map.hpp:749 */

typedef struct kh_table kh_table;

const char *kh_last_error(void);
/* number of CUDA devices visible (<= 0: none) */
int kh_device_count(void);
/* number of kernels this library launched since load / since the last reset (bench.py's gpu_launches) */
uint64_t kh_launch_count(void);
void kh_reset_launch_count(void);

/* ---- pinned host memory helpers (for the *_host fast path) ---------------------------- */
void *kh_host_alloc(size_t bytes);   /* cudaHostAlloc; NULL on failure */
void kh_host_free(void *p);

/* ---- table lifetime and sizing -------------------------------------------------------
 * replaces: MAP() default ctor, set_max/min_load_factor (KH/robinhood_offset_hashmap_ptr.hpp:407-425,
 * 631-640), reserve/rehash (:896-981), capacity/size/clear. */
kh_table *kh_create(int device, int map_kind, int key_words, int hash_kind, uint32_t seed, int reducer);
void kh_destroy(kh_table *t);
int kh_set_load_factors(kh_table *t, double min_load, double max_load);
/* accepted and ignored (software prefetch distances have no meaning on the GPU);
 * replaces set_insert_lookahead / set_query_lookahead (:646-653) */
int kh_set_lookahead(kh_table *t, int insert_lookahead, int query_lookahead);
/* capacity := next_pow2(ceil(n / max_load)) if that is larger than the current one (bulk GPU rebuild) */
int kh_reserve(kh_table *t, uint64_t n);
/* force capacity to next_pow2(buckets) (>= what size() needs); replaces rehash(b) */
int kh_rehash(kh_table *t, uint64_t slots);
uint64_t kh_capacity(const kh_table *t);
uint64_t kh_size(const kh_table *t);
double kh_load_factor(const kh_table *t);
int kh_clear(kh_table *t);
int kh_device(const kh_table *t);
int kh_key_words(const kh_table *t);
/* the CUDA stream the *_dev entry points of this table launch on (a cudaStream_t), for event timing */
void *kh_stream(kh_table *t);
/* CUDA-event timing of the probe kernels on their launching stream: enable, run operations, then read
 * the accumulated milliseconds and launch count of which = 0 (table insert kernels: tile insert + global insert kernel) /
 * 1 (lookup kernels) / 2 (level-1 scatter of the bulk insert) / 3 (stand-alone HyperLogLog pass) / 4 (every kernel of a
 * rebuild: reserve, rehash, shrink after erase) / 5 (level-2 split of the bulk insert) */
int kh_kernel_timing(kh_table *t, int enable);
int kh_kernel_time(kh_table *t, int which, double *sum_ms, uint64_t *launches);
/* probe geometry (experiments; empty table only): lane_bytes 16|32|64, lanes_per_bucket 1|2|4|8 (64: 1 only) -- see DESIGN.md */
int kh_set_geometry(kh_table *t, int lane_bytes, int lanes_per_bucket);

/* L2-windowed bulk insert: COUNT batches of >= min_keys keys into a table of
 * >= 4 * window_bytes are first sorted by table region (window_bytes each, power of two) so the probe
 * kernel works inside the L2.  Default: on, 32 MiB, 4 Mi keys.  Results do not depend on it. */
int kh_set_insert_window(kh_table *t, int enable, uint64_t window_bytes, uint64_t min_keys);

/* HyperLogLog distinct estimate of a key batch over the table's store hash, precision 12
 * (replaces get_hll().update(keys,n) + estimate(), KH/hyperloglog64.hpp:176-239; driver
 * "estimate" phase benchmark/BenchmarkHashTables.cpp:1232-1248).  Registers are bit-identical
 * to the reference's; `regs_out` (4096 bytes, may be NULL) receives them. */
double kh_estimate_distinct_host(kh_table *t, const uint64_t *keys, uint64_t n, uint8_t *regs_out);
double kh_estimate_distinct_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, uint8_t *regs_out);

/* ---- batch insert --------------------------------------------------------------------
 * replaces insert(vector<Key> const&, T default) / insert(IT,IT) / insert_no_estimate
 * (KH/robinhood_offset_hashmap_ptr.hpp:2783-2885) and the per-key insert loops of RH/LP
 * (KH/hashmap_robinhood.hpp:522-717, KH/hashmap_linearprobe.hpp:430-573).
 * vals == NULL: every key carries `default_val`.  estimate != 0: HLL-presize first (insert());
 * 0: grow by doubling only (insert_no_estimate()).  Returns the number of NEW keys, < 0 on error. */
int64_t kh_insert_host(kh_table *t, const uint64_t *keys, const uint32_t *vals, uint64_t n,
                       uint32_t default_val, int estimate);
int64_t kh_insert_dev(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n,
                      uint32_t default_val, int estimate);

/* Same, for an array of std::pair<Kmer, T> records as the reference's insert(IT, IT) /
 * insert(std::vector<value_type> const&) take them (:2783-2796, :2802-2885): `stride_bytes` =
 * sizeof(value_type) (multiple of 8), key words at offset 0, the T value (`val_bytes` = 1, 2 or 4) at
 * `val_offset`.  The records are copied as they are and de-interleaved by a kernel. */
int64_t kh_insert_pairs_host(kh_table *t, const void *pairs, uint64_t n, uint32_t stride_bytes,
                             uint32_t val_offset, uint32_t val_bytes, int estimate);

/* ---- batch queries -------------------------------------------------------------------
 * kh_count_*: replaces count(out, b, e) / exists (:3251-3328): out[i] = 0/1; returns #present.
 * kh_find_*:  replaces find<OT>(out, b, e, nonexistent) (:3376-3417): out[i] = value or `nonexistent`;
 *             returns #found.
 * kh_find_existing_*: replaces find_existing (:3425-3454): hits only, in query order; returns #hits. */
int64_t kh_count_host(kh_table *t, const uint64_t *keys, uint64_t n, uint8_t *out);
int64_t kh_count_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, uint8_t *d_out);
int64_t kh_find_host(kh_table *t, const uint64_t *keys, uint64_t n, uint32_t *out, uint32_t nonexistent);
int64_t kh_find_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, uint32_t *d_out, uint32_t nonexistent);
int64_t kh_find_existing_host(kh_table *t, const uint64_t *keys, uint64_t n, uint64_t *out_keys, uint32_t *out_vals);
int64_t kh_find_existing_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, uint64_t *d_out_keys, uint32_t *d_out_vals);

/* replaces update(...) (:3467-3495; KH/hashmap_robinhood.hpp:1274): overwrite the value of keys
 * that are present, never inserts; returns #updated. */
int64_t kh_update_host(kh_table *t, const uint64_t *keys, const uint32_t *vals, uint64_t n);
int64_t kh_update_dev(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n);

/* ---- batch erase ---------------------------------------------------------------------
 * replaces erase(Key const*, Key const*) (:3768-3790; RH :1421-1440; LP :1032-1051):
 * tombstone-free; returns #keys removed (a key repeated in the batch counts once); shrinks
 * (bulk rebuild) when the load drops below min_load, like the reference. */
int64_t kh_erase_host(kh_table *t, const uint64_t *keys, uint64_t n);
int64_t kh_erase_dev(kh_table *t, const uint64_t *d_keys, uint64_t n);

/* ---- export --------------------------------------------------------------------------
 * replaces to_vector() / keys() (:861-878): unordered; writes at most `cap` entries, returns size(). */
int64_t kh_export_host(kh_table *t, uint64_t *out_keys, uint32_t *out_vals, uint64_t cap);
int64_t kh_export_dev(kh_table *t, uint64_t *d_out_keys, uint32_t *d_out_vals, uint64_t cap);

/* ---- batched hashing (parity surface for the hash functors) --------------------------
 * replaces Hash::operator()(Key const*, n, out*) (KH/hash_new.hpp:390-1134 batch mode).
 * out[i] is the functor's value zero-extended to 64 bits. */
int kh_hash_host(int device, int hash_kind, uint32_t seed, int key_words, const uint64_t *keys, uint64_t n, uint64_t *out);
int kh_hash_dev(int device, int hash_kind, uint32_t seed, int key_words, const uint64_t *d_keys, uint64_t n, uint64_t *d_out);

/* ---- k-mer generation ----------------------------------------------------------------
 * replaces KmerParser::operator() over NSplitSequencesIterator + lex_less
 * (KI/io/kmer_parser.hpp:232-293, KI/io/filtered_sequence_iterator.hpp:225-247,429-440,
 * KI/common/kmer.hpp:716-726,1078-1087, KI/common/kmer_transform.hpp:109-125):
 * read r is bases[offsets[r] .. offsets[r+1]); '\n' '\r' skipped; 'N'/'n' split; other bytes
 * through the DNA table (non-ACGT -> A).  Output order = read order, then position.
 * out_keys == NULL: count only.  Returns the number of k-mers. */
int64_t kh_kmers_from_reads_host(int device, const char *bases, const uint64_t *offsets, uint64_t n_reads,
                                 int k, int canonical, uint64_t *out_keys, uint64_t out_cap);
int64_t kh_kmers_from_reads_dev(int device, const char *d_bases, const uint64_t *d_offsets, uint64_t n_reads,
                                int k, int canonical, uint64_t *d_out_keys, uint64_t out_cap);
/* replaces transform_input with lex_less (KI/containers/distributed_map_base.hpp:287-301): in-place allowed */
int kh_canonicalize_dev(int device, int k, const uint64_t *d_in, uint64_t n, uint64_t *d_out);
/* Other alphabets (the reference instantiates testKmerIndex 31 x {DNA, DNA5, DNA16}, benchmark/CMakeLists.txt:175-196):
 * alphabet = 4 (DNA, 2 bits), 5 (DNA5 = the 3-bit DNA6 table, KI/common/alphabets.hpp:197-292,746) or 16 (DNA16, 4-bit IUPAC codes,
 * :648-739); keys of up to two words, i.e. k <= 64 / 42 / 32.  Reverse complement of DNA5 / DNA16 = reversal of all bits of the value
 * (KI/common/kmer.hpp:1476-1497).  The table, hashes and exchange take these keys like any 8- or 16-byte key. */
int kh_kmer_words_alpha(int alphabet, int k);
int64_t kh_kmers_from_reads_alpha_dev(int device, int alphabet, const char *d_bases, const uint64_t *d_offsets, uint64_t n_reads,
                                      int k, int canonical, uint64_t *d_out_keys, uint64_t out_cap);
int64_t kh_kmers_from_reads_alpha_host(int device, int alphabet, const char *bases, const uint64_t *offsets, uint64_t n_reads,
                                       int k, int canonical, uint64_t *out_keys, uint64_t out_cap);
int kh_canonicalize_alpha_host(int device, int alphabet, int k, const uint64_t *in, uint64_t n, uint64_t *out);
int kh_revcomp_alpha_host(int device, int alphabet, int k, const uint64_t *in, uint64_t n, uint64_t *out);
int kh_canonicalize_host(int device, int k, const uint64_t *in, uint64_t n, uint64_t *out);
int kh_revcomp_host(int device, int k, const uint64_t *in, uint64_t n, uint64_t *out);

/* ---- hash partitioning for the multi-GPU exchange ------------------------------------
 * replaces assign_count_permute + permute_by_bucketid
 * (KH/distributed_batched_robinhood_map.hpp:164-178,291-380,632-741): rank =
 * DistHash_{seed}(key) & (p-1) for power-of-two p, % p otherwise (in the functor's result
 * width); STABLE scatter so each destination's keys are contiguous and in input order.
 * d_out_keys: n keys permuted; d_out_perm (may be NULL): original index of every permuted
 * position; counts (HOST, p entries): keys per destination. */
int kh_partition_dev(int device, int hash_kind, uint32_t seed, int key_words, uint32_t p,
                     const uint64_t *d_keys, uint64_t n, uint64_t *d_out_keys, uint32_t *d_out_perm,
                     uint64_t *counts);
/* Same partition for callers that do not need the reference's STABLE order (the insert path: COUNT is
 * commutative; the query path: d_out_perm restores input order): each destination's keys are contiguous,
 * order inside a destination unspecified.  One histogram pass + one tile-sorting scatter pass
 * (16 B/key of streaming traffic), p <= 1024. */
int kh_partition_fast_dev(int device, int hash_kind, uint32_t seed, int key_words, uint32_t p,
                          const uint64_t *d_keys, uint64_t n, uint64_t *d_out_keys, uint32_t *d_out_perm,
                          uint64_t *counts);
/* kh_partition_fast_dev without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, same results as kh_partition_fast_dev.  The kernels' scratch (histogram, cursors, totals: 3 counters per
 * destination) lies in a static device array of KH_PARTITION_NOALLOC_MAX_P * 24 bytes that every GPU loading the library
 * holds until exit; p above KH_PARTITION_NOALLOC_MAX_P is rejected with KH_ERR_INVALID.  The call runs on the calling
 * thread's per-thread stream and returns when the results are complete; it calls no cudaMalloc / cudaFree / pool routine.
 * All calls on one device share the array: calls that overlap on the GPU (several host threads) race on it. */
#define KH_PARTITION_NOALLOC_MAX_P 1024
int kh_partition_fast_noalloc_dev(int device, int hash_kind, uint32_t seed, int key_words, uint32_t p,
                                  const uint64_t *d_keys, uint64_t n, uint64_t *d_out_keys, uint32_t *d_out_perm,
                                  uint64_t *counts);
int kh_rank_host(int device, int hash_kind, uint32_t seed, int key_words, uint32_t p,
                 const uint64_t *keys, uint64_t n, uint32_t *out_rank);
/* out[perm[i]] = in[i] for 1- or 4-byte elements (results of a permuted query back to input order) */
int kh_unpermute_dev(int device, int elem_bytes, const void *d_in, const uint32_t *d_perm, uint64_t n, void *d_out);

/* ---- one rank of a multi-process (one process per GPU) distributed insert ---------------
 * replaces assign_count_estimate_permute (KH/distributed_batched_robinhood_map.hpp:398-625: rank per key,
 * per-rank counts, PER-DESTINATION HyperLogLog registers, permute) and the receiving half of insert_p
 * (:910-1179: merge the estimates, reserve, insert).  `t` is this rank's table (store hash, key width).
 * Sender:   khx_plan_scatter_dev partitions d_keys by owner (p <= 8) into d_out_keys (destination segments
 *           contiguous, order inside unspecified), fills counts[p] (HOST) and d_meta = p x khx_meta_words()
 *           uint64 (DEVICE): for every destination its key count, its keys' histogram over table regions and
 *           their HLL registers.  Keys and metadata travel by all-to-all (NCCL; not done here).
 * Receiver: khx_insert_recv_dev takes the n received keys and the p received metadata blocks, merges them
 *           (registers: max, histograms: sum), presizes the table from the merged estimate and inserts with
 *           value default_val under the table's reducer (COUNT).  Returns #new keys. */
int khx_meta_words(void);
int khx_plan_scatter_dev(kh_table *t, int dist_hash, uint32_t dist_seed, uint32_t p, const uint64_t *d_keys, uint64_t n,
                         uint64_t *d_meta, uint64_t *d_out_keys, uint64_t *counts);
int64_t khx_insert_recv_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, const uint64_t *d_recv_meta, uint32_t p,
                            uint32_t default_val);

/* Fused partition + exchange over peer memory (NVLink): instead of permuting into a send buffer and calling an
 * all-to-all, the rank scatter stores every key straight into the receive buffer of its owner GPU.  The
 * receive buffers are cudaMalloc'ed per rank and mapped into the other processes of the box with CUDA IPC
 * (khx_peer_handle -> exchange the 64 bytes -> khx_peer_open); khx_peer_set maps a plain device pointer
 * (same-process peers).  khx_plan_dev = the metadata pass of khx_plan_scatter_dev alone.  dst_offsets[r] /
 * dst_counts[r] (HOST, p entries): where this rank's segment starts in rank r's buffer and how long it is
 * (from the p x p count matrix every rank gathers).  Ordering between ranks is the caller's (see distributed.py). */
typedef struct khx_peer khx_peer;
khx_peer *khx_peer_create(int device, uint32_t rank, uint32_t p, uint64_t capacity_keys, int key_words);
void khx_peer_destroy(khx_peer *x);
uint64_t khx_peer_capacity(const khx_peer *x);
uint64_t *khx_peer_recv_ptr(khx_peer *x);
int khx_peer_handle(khx_peer *x, void *out64);
int khx_peer_open(khx_peer *x, uint32_t r, const void *handle64);
int khx_peer_set(khx_peer *x, uint32_t r, uint64_t *d_ptr);
int khx_plan_dev(kh_table *t, int dist_hash, uint32_t dist_seed, uint32_t p, const uint64_t *d_keys, uint64_t n,
                 uint64_t *d_meta);
int khx_scatter_peers_dev(kh_table *t, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n,
                          khx_peer *x, const uint64_t *dst_offsets, const uint64_t *dst_counts);

/* Single-pass sender (assign_count_estimate_permute :398-625 and the exchange as ONE kernel): every receive buffer is cut
 * into 2 halves x p slabs of slab_cap keys; this rank stores its keys for owner r into slab `rank` of half `parity` of
 * r's buffer while the same pass fills d_meta (counts, region histograms, HLL registers per owner) -- no offset depends
 * on another rank, so nothing has to be gathered before the stores.  A slab that would overflow is cut off; its true
 * count is in d_meta[r][0], so after the metadata all-gather every rank sees the overflow and the caller redoes the batch
 * with khx_plan_dev + khx_scatter_peers_dev.  The halves alternate between batches (the all-gather of batch i+1 orders
 * batch i+2 behind every rank's reads of batch i).  khx_insert_recv_slabs_dev = khx_insert_recv_dev reading the p slabs
 * of this rank's own buffer in place (counts[g], HOST: keys sender g stored). */
int khx_plan_scatter_slabs_dev(kh_table *t, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n, khx_peer *x,
                               uint64_t slab_cap, int parity, uint64_t *d_meta);
int64_t khx_insert_recv_slabs_dev(kh_table *t, khx_peer *x, uint64_t slab_cap, int parity, const uint64_t *counts,
                                  const uint64_t *d_recv_meta, uint32_t default_val);

/* Query exchange over the slabs: replaces the exchange of count_p / find_p / erase_p
 * (KH/distributed_batched_robinhood_map.hpp:1258-1516,1619,2169) and the reply of ialltoallv_and_query_one_to_one
 * (KH/incremental_mxx.hpp:4403).  Buffers come from khx_peer_create_q: p slabs of slab_cap keys followed by a reply region
 * (4 bytes per key).  Asker: khx_query_scatter_slabs_dev stores its keys for owner r into slab `rank` of r's buffer,
 * d_perm[r * slab_cap + i] (DEVICE, p * slab_cap u32) = input index of the i-th key sent to r, d_counts[r] (DEVICE) = keys
 * for r (a slab that would overflow is cut off, the count stays true: the caller sees it in the gathered counts and falls
 * back).  Owner, once every asker's stores have landed: khx_query_slabs_dev looks the slabs up (op 0 find -> u32, 1 count ->
 * u8, 2 erase) and stores the answers straight into slab `rank` of asker g's reply region; returns #hits / #erased.
 * Asker, once every owner's answers have landed: khx_unpermute_slabs_dev puts them into input order in d_out. */
khx_peer *khx_peer_create_q(int device, uint32_t rank, uint32_t p, uint64_t capacity_keys, int key_words);
/* Barrier between the ranks over peer memory (flags in a header behind every receive buffer; replaces the MPI_Barrier /
 * MPI_Waitall points of khmxx, KH/incremental_mxx.hpp:3437-4403): returns once every rank has called it with the same
 * epoch (1, 2, 3, ... per buffer set); all peer stores a rank issued before its call are then visible to their owners.
 * d_counts_row (DEVICE, p words or NULL) is published to every rank, counts_matrix (HOST, p x p or NULL) receives the
 * matrix [sender][owner].  KH_ERR_CUDA when a peer does not arrive within 10 s. */
int khx_peer_sync_dev(kh_table *t, khx_peer *x, uint64_t epoch, const uint64_t *d_counts_row, uint64_t *counts_matrix);
/* Insert exchange without NCCL: after khx_plan_scatter_slabs_dev, push metadata block r of d_meta into owner r's metadata
 * area (half `parity`, slot of this rank) by peer stores and pass the barrier, which also delivers the count matrix
 * (d_meta[r][0]).  khx_peer_meta_ptr = this rank's area of that half: p blocks of khx_meta_words() words, block g from
 * sender g -- the d_recv_meta of khx_insert_recv_slabs_dev. */
int khx_push_meta_sync_dev(kh_table *t, khx_peer *x, int parity, const uint64_t *d_meta, uint64_t epoch, uint64_t *counts_matrix);
/* The insert sender without metadata: dist hash -> owner -> store into slab `rank` of half `parity` of the owner's buffer.
 * d_counts (DEVICE, p words) = this rank's row of the count matrix, to be handed to khx_peer_sync_dev; reset_counts = 0
 * continues a batch that arrives in chunks (e.g. behind a PCIe copy): the counts accumulate.  Asynchronous on the table's stream.  khx_insert_recv_slabs_dev accepts d_recv_meta == NULL (the receiver's level-1 pass over the slabs
 * computes the HyperLogLog registers the reference ships with the counts, KH/distributed_batched_robinhood_map.hpp:1024-1060). */
int khx_scatter_slabs_dev(kh_table *t, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n, khx_peer *x,
                          uint64_t slab_cap, int parity, uint64_t *d_counts, int reset_counts);
/* The exchange overlapped with the local insertion (the reference's ialltoallv_and_modify, KH/incremental_mxx.hpp:3437-3645):
 * the batch goes in `chunks` pieces; per piece a sender kernel, a flag barrier and -- on a second stream, launched before
 * the host knows any count -- the level-1 pass of the bulk insert over the p sub-slabs that piece filled, so that the
 * receiver works on piece c while the senders store piece c + 1.  slab_cap must be a multiple of 2 * chunks.  Consumes the
 * epochs epoch0 .. epoch0 + chunks - 1.  counts_matrix (HOST, p x p) = keys sender g had for owner r.  *overflowed = 1: a
 * sub-slab was too small (the same verdict on every rank); the table is untouched, repeat the batch on the exact path.
 * COUNT tables on the bulk route only (khx_insert_recv_slabs_dev serves the rest).  Returns the number of new keys. */
int64_t khx_insert_pipelined_dev(kh_table *t, khx_peer *x, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n,
                                 uint64_t slab_cap, int parity, int chunks, uint64_t epoch0, uint32_t default_val,
                                 uint64_t *counts_matrix, int *overflowed);
/* The same schedule with the COPY ENGINES on the links (what mxx's non-blocking alltoallv leaves to the MPI progress engine,
 * KH/incremental_mxx.hpp:3437-3645): every piece is partitioned by owner into a staging buffer in this GPU's memory (its own
 * keys straight into its own slab), cudaMemcpyAsync moves each (piece, owner) run into the owner's slab on a copy stream, the
 * flag barrier of the piece follows on that stream, and the SMs meanwhile partition the next piece and run the level-1 pass
 * of the pieces that have arrived.  Arguments, results and limits as khx_insert_pipelined_dev. */
int64_t khx_insert_ce_dev(kh_table *t, khx_peer *x, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n,
                          uint64_t slab_cap, int parity, int chunks, uint64_t epoch0, uint32_t default_val,
                          uint64_t *counts_matrix, int *overflowed);
/* k-mer generation fused with the sender (SURVEY.md 2.4 "fused with k-mer gen on the way in"; KmerParser -> transform_input ->
 * assign_count_permute of the reference, KI/io/kmer_parser.hpp:232-293, KH/distributed_batched_robinhood_map.hpp:632-741): reads ->
 * canonical k-mers -> owner -> peer stores, one kernel, the k-mers never written to this GPU's memory.  d_counts as for
 * khx_scatter_slabs_dev.  KH_ERR_FULL: EOL-dense input the tile kernel cannot decide -- generate with kh_kmers_from_reads_dev and
 * send with khx_scatter_slabs_dev instead (what was stored meanwhile is to be ignored). */
int khx_kmers_scatter_slabs_dev(kh_table *t, const char *d_bases, const uint64_t *d_offsets, uint64_t n_reads, int k, int canonical,
                                int dist_hash, uint32_t dist_seed, khx_peer *x, uint64_t slab_cap, int parity, uint64_t *d_counts);
/* Teardown of IPC-mapped buffers is collective: every rank khx_peer_close_imports -> a barrier of the caller's -> every
 * rank khx_peer_destroy (which frees the exported allocation).  khx_peer_poisoned: a barrier on these buffers timed out
 * (on this rank or, seen through the poison flag, on a peer); every later khx_peer_sync_dev fails until they are recreated. */
void khx_peer_close_imports(khx_peer *x);
int khx_peer_poisoned(const khx_peer *x);
uint64_t *khx_peer_meta_ptr(khx_peer *x, int parity);
int khx_query_scatter_slabs_dev(kh_table *t, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n, khx_peer *x,
                                uint64_t slab_cap, uint32_t *d_perm, uint64_t *d_counts);
int64_t khx_query_slabs_dev(kh_table *t, khx_peer *x, int op, uint64_t slab_cap, const uint64_t *counts, uint32_t nonexistent);
int khx_unpermute_slabs_dev(kh_table *t, khx_peer *x, int elem_bytes, uint64_t slab_cap, const uint32_t *d_perm,
                            const uint64_t *d_counts, uint64_t n, void *d_out);

/* ---- single-process multi-GPU map ------------------------------------------------------
 * replaces dsc::counting_batched_robinhood_map / batched_robinhood_map
 * (KH/distributed_batched_robinhood_map.hpp:748-832,1187,1535,1901,2428,2542-2649): one table per
 * GPU, owner = DistHash_{9876543}(canon(key)) mod p, exchange by direct peer stores over NVLink
 * inside the partition kernel.  Results come back in INPUT order. */
typedef struct khd_map khd_map;
khd_map *khd_create(int n_gpus, int key_words, int store_hash, int dist_hash, int reducer, int canonical_k);
void khd_destroy(khd_map *m);
int khd_n_gpus(const khd_map *m);
/* 1 when partition+exchange run as one kernel storing into peer GPUs' buffers (peer access available) */
int khd_uses_peer_stores(const khd_map *m);
int64_t khd_insert_host(khd_map *m, const uint64_t *keys, uint64_t n, uint32_t default_val);
/* (key, value) pairs for FIRST / LAST maps (benchmark/BenchmarkDistHashTables.cpp:976-1020): the owner sees the pairs sender by
 * sender in rank order, input order inside a sender, as the reference's owner rank does after its alltoallv */
int64_t khd_insert_pairs_host(khd_map *m, const uint64_t *keys, const uint32_t *vals, uint64_t n);
int khd_set_load_factors(khd_map *m, double min_load, double max_load);
int64_t khd_count_host(khd_map *m, const uint64_t *keys, uint64_t n, uint8_t *out);
int64_t khd_find_host(khd_map *m, const uint64_t *keys, uint64_t n, uint32_t *out, uint32_t nonexistent);
int64_t khd_erase_host(khd_map *m, const uint64_t *keys, uint64_t n);
uint64_t khd_size(const khd_map *m);
uint64_t khd_local_size(const khd_map *m, int rank);
int64_t khd_export_rank_host(khd_map *m, int rank, uint64_t *out_keys, uint32_t *out_vals, uint64_t cap);

/* ---- roofline probes (measurement support, not part of the reference API) --------------
 * Touch `n_access` uniformly random `bytes_per_access`-byte aligned pieces of a `buf_bytes`
 * buffer: mode 0 = read, 1 = read + atomicAdd (RMW).  Returns the kernel's milliseconds. */
double kh_probe_random_access(int device, uint64_t buf_bytes, uint64_t n_access, int bytes_per_access, int mode, int iters);
double kh_probe_stream_copy(int device, uint64_t bytes, int iters);

#ifdef __cplusplus
}
#endif
#endif /* KMERHASH_B200_H_ */
