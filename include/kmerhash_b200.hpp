// kmerhash_b200.hpp -- header-only C++ facade over the C ABI (kmerhash_b200.h).
//
// The classes carry the names and member signatures of the reference containers so that a driver
// written against the reference (benchmark/BenchmarkHashTables.cpp:588-1364) compiles against
// either with a namespace switch:
//   ::fsc::hashmap_robinhood_offsets_reduction / _offsets / _offsets_count
//        include/kmerhash/robinhood_offset_hashmap_ptr.hpp:196-202, 4094, 4099
//   ::fsc::hashmap_robinhood_doubling      include/kmerhash/hashmap_robinhood.hpp:126
//   ::fsc::hashmap_linearprobe_doubling    include/kmerhash/hashmap_linearprobe.hpp:98
// Differences that follow from the data living in HBM: there are no iterators into the table
// (begin()/end() are replaced by to_vector()/keys()), the Hash template argument is a TAG naming
// one of the reference's functors (hash values are bit-identical), and Equal/Allocator arguments
// are accepted and ignored (equality is whole-word equality, as KI/common/kmer.hpp:775-779).
// Errors surface as the exception types the reference throws.
#ifndef KMERHASH_B200_HPP_
#define KMERHASH_B200_HPP_

#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "kmerhash_b200.h"

namespace khb {

// ---- key type: layout of bliss::common::Kmer<K, DNA, uint64_t> (KI/common/kmer.hpp:115-170) ----
template <unsigned K>
struct Kmer {
  static constexpr unsigned size = K;
  static constexpr unsigned nWords = (2 * K + 63) / 64;
  static_assert(nWords == 1 || nWords == 2, "k <= 64 (2-bit DNA) on this path");
  uint64_t data[nWords];
  Kmer() { for (unsigned i = 0; i < nWords; ++i) data[i] = 0; }
  explicit Kmer(uint64_t w0, uint64_t w1 = 0) { data[0] = w0; if (nWords > 1) data[nWords - 1] = w1; sanitize(); }
  void sanitize() {   // do_sanitize, kmer.hpp:1286: unused top bits forced to 0
    const unsigned used = 2 * K - 64 * (nWords - 1);
    if (used < 64) data[nWords - 1] &= (~0ULL) >> (64 - used);
  }
  bool operator==(Kmer const &o) const { return std::memcmp(data, o.data, sizeof(data)) == 0; }
  bool operator!=(Kmer const &o) const { return !(*this == o); }
  bool operator<(Kmer const &o) const {   // most significant word first, kmer.hpp:790-845
    for (int i = (int)nWords - 1; i >= 0; --i) if (data[i] != o.data[i]) return data[i] < o.data[i];
    return false;
  }
};

// ---- hash tags = the reference's StoreHash choices (benchmark/BenchmarkHashTables.cpp:131-166) ----
namespace hash {
template <int KIND, uint32_t SEED = KH_DEFAULT_SEED> struct tag { static constexpr int kind = KIND; static constexpr uint32_t seed = SEED; };
using identity     = tag<KH_HASH_IDEN>;
using murmur       = tag<KH_HASH_MURMUR64>;
using murmur_x86   = tag<KH_HASH_MURMUR_X86>;
using murmur3avx64 = tag<KH_HASH_MURMUR64AVX>;
using murmur32     = tag<KH_HASH_MURMUR32>;
using murmur3avx32 = tag<KH_HASH_MURMUR32AVX>;
using murmur3sse32 = tag<KH_HASH_MURMUR32SSE>;
using farm         = tag<KH_HASH_FARM>;
using farm32       = tag<KH_HASH_FARM32>;
using clhash       = tag<KH_HASH_CLHASH>;
using crc32c       = tag<KH_HASH_CRC32C, KH_DEFAULT_CRC_SEED>;
}  // namespace hash

// ---- reducers (include/kmerhash/robinhood_offset_hashmap_ptr.hpp:85-97) ----
struct DiscardReducer { static constexpr int kind = KH_REDUCE_FIRST; };
struct ReplaceReducer { static constexpr int kind = KH_REDUCE_LAST; };
struct PlusReducer    { static constexpr int kind = KH_REDUCE_COUNT; };   // ::std::plus<T>
template <typename R> struct reducer_kind { static constexpr int value = R::kind; };
template <typename T> struct reducer_kind<std::plus<T>> { static constexpr int value = KH_REDUCE_COUNT; };

namespace detail {

[[noreturn]] inline void raise(int64_t code, const char *what) {
  std::string msg = std::string(what) + ": " + kh_last_error();
  switch (code) {
    case KH_ERR_INVALID: throw std::invalid_argument(msg);
    case KH_ERR_NOMEM:   throw std::length_error(msg);     // mem_utils.hpp:47
    default:             throw std::logic_error(msg);      // hashmap_linearprobe.hpp:503
  }
}
inline int64_t ok(int64_t rc, const char *what) { if (rc < 0) raise(rc, what); return rc; }

// engine shared by the three container families
template <typename Key, typename T, typename HashTag, int REDUCER, int MAPKIND>
class table_base {
  static_assert(std::is_trivially_copyable<Key>::value && (sizeof(Key) == 8 || sizeof(Key) == 16),
                "Key must be an 8- or 16-byte POD (Kmer<K<=32> / Kmer<K<=64> / uint64_t)");
  static_assert(std::is_integral<T>::value && sizeof(T) <= 4, "mapped type: integral, at most 32 bits");

 public:
  using key_type = Key;
  using mapped_type = T;
  using value_type = std::pair<Key, T>;
  using size_type = size_t;
  static constexpr int key_words = sizeof(Key) / 8;

  explicit table_base(size_t initial_capacity = 128, int device = 0) {
    h_ = kh_create(device, MAPKIND, key_words, HashTag::kind, HashTag::seed, REDUCER);
    if (!h_) raise(kh_device_count() <= 0 ? KH_ERR_CUDA : KH_ERR_INVALID, "kh_create");
    if (initial_capacity > 128) rehash(initial_capacity);
  }
  ~table_base() { if (h_) kh_destroy(h_); }
  table_base(table_base const &) = delete;
  table_base &operator=(table_base const &) = delete;
  table_base(table_base &&o) noexcept : h_(o.h_), min_(o.min_), max_(o.max_) { o.h_ = nullptr; }

  // ---- sizing ----
  void set_max_load_factor(double v) { max_ = v; ok(kh_set_load_factors(h_, min_, max_), "set_max_load_factor"); }
  void set_min_load_factor(double v) { min_ = v; ok(kh_set_load_factors(h_, min_, max_), "set_min_load_factor"); }
  double get_max_load_factor() const { return max_; }
  double get_min_load_factor() const { return min_; }
  double get_load_factor() const { return kh_load_factor(h_); }
  void set_insert_lookahead(unsigned char v) { kh_set_lookahead(h_, v, 0); }   // no meaning on the GPU
  void set_query_lookahead(unsigned char v) { kh_set_lookahead(h_, 0, v); }
  void reserve(size_t n) { ok(kh_reserve(h_, n), "reserve"); }
  void rehash(size_t buckets) { ok(kh_rehash(h_, buckets), "rehash"); }
  size_t capacity() const { return kh_capacity(h_); }
  size_t size() const { return kh_size(h_); }
  bool empty() const { return size() == 0; }
  void clear() { ok(kh_clear(h_), "clear"); }
  kh_table *native_handle() { return h_; }
  // get_hll().update(keys, n) + estimate()  (hyperloglog64.hpp:176-239)
  double estimate(Key const *keys, size_t n) {
    double e = kh_estimate_distinct_host(h_, reinterpret_cast<const uint64_t *>(keys), n, nullptr);
    if (e < 0) raise(KH_ERR_CUDA, "estimate");
    return e;
  }

  // ---- export ----
  std::vector<value_type> to_vector() const {
    const size_t n = size();
    std::vector<Key> k(n); std::vector<uint32_t> v(n);
    ok(kh_export_host(h_, reinterpret_cast<uint64_t *>(k.data()), v.data(), n), "to_vector");
    std::vector<value_type> out(n);
    for (size_t i = 0; i < n; ++i) out[i] = value_type(k[i], static_cast<T>(v[i]));
    return out;
  }
  std::vector<Key> keys() const {
    const size_t n = size();
    std::vector<Key> k(n);
    ok(kh_export_host(h_, reinterpret_cast<uint64_t *>(k.data()), nullptr, n), "keys");
    return k;
  }

 protected:
  static const uint64_t *kp(Key const *p) { return reinterpret_cast<const uint64_t *>(p); }
  static uint32_t val_offset() {
    value_type v{};
    return static_cast<uint32_t>(reinterpret_cast<const char *>(&v.second) - reinterpret_cast<const char *>(&v));
  }
  size_t insert_pairs(value_type const *b, value_type const *e, bool estimate) {
    return (size_t)ok(kh_insert_pairs_host(h_, b, (uint64_t)(e - b), (uint32_t)sizeof(value_type), val_offset(),
                                           (uint32_t)sizeof(T), estimate ? 1 : 0), "insert");
  }
  size_t insert_keys(Key const *b, Key const *e, T const &dflt, bool estimate) {
    return (size_t)ok(kh_insert_host(h_, kp(b), nullptr, (uint64_t)(e - b), (uint32_t)dflt, estimate ? 1 : 0), "insert");
  }
  size_t lookup_vals(Key const *b, Key const *e, uint32_t *out, uint32_t nonexistent) const {
    return (size_t)ok(kh_find_host(h_, kp(b), (uint64_t)(e - b), out, nonexistent), "find");
  }
  size_t lookup_flags(Key const *b, Key const *e, uint8_t *out) const {
    return (size_t)ok(kh_count_host(h_, kp(b), (uint64_t)(e - b), out), "count");
  }
  std::vector<value_type> hits(Key const *b, Key const *e) const {
    const size_t n = e - b;
    std::vector<Key> k(n); std::vector<uint32_t> v(n);
    const size_t m = (size_t)ok(kh_find_existing_host(h_, kp(b), n, reinterpret_cast<uint64_t *>(k.data()), v.data()), "find_existing");
    std::vector<value_type> out(m);
    for (size_t i = 0; i < m; ++i) out[i] = value_type(k[i], static_cast<T>(v[i]));
    return out;
  }
  size_t erase_keys(Key const *b, Key const *e) { return (size_t)ok(kh_erase_host(h_, kp(b), (uint64_t)(e - b)), "erase"); }
  size_t update_pairs(value_type const *b, value_type const *e) {
    const size_t n = e - b;
    std::vector<Key> k(n); std::vector<uint32_t> v(n);
    for (size_t i = 0; i < n; ++i) { k[i] = b[i].first; v[i] = (uint32_t)b[i].second; }
    return (size_t)ok(kh_update_host(h_, kp(k.data()), v.data(), n), "update");
  }

  kh_table *h_ = nullptr;
  double min_ = MAPKIND == KH_MAP_ROBINHOOD ? 0.4 : 0.2;
  double max_ = MAPKIND == KH_MAP_ROBINHOOD ? 0.9 : (MAPKIND == KH_MAP_LINEARPROBE ? 0.6 : 0.8);
};

}  // namespace detail

// =====================================================================================================
// hashmap_robinhood_offsets_reduction  (include/kmerhash/robinhood_offset_hashmap_ptr.hpp:196-202)
// Batch interface of the reference: insert :2783-2885, count :3311-3328, find :3376-3417,
// find_existing :3425-3454, update :3467-3495, erase :3768-3790.
// =====================================================================================================
template <typename Key, typename T, typename HashTag = hash::murmur3avx64, typename Equal = std::equal_to<Key>,
          typename Reducer = DiscardReducer, typename Allocator = std::allocator<std::pair<const Key, T>>>
class hashmap_robinhood_offsets_reduction
    : public detail::table_base<Key, T, HashTag, reducer_kind<Reducer>::value, KH_MAP_ROBINHOOD_OFFSETS> {
  using base = detail::table_base<Key, T, HashTag, reducer_kind<Reducer>::value, KH_MAP_ROBINHOOD_OFFSETS>;

 public:
  using typename base::value_type;
  using base::base;

  // insert(): HLL estimate -> reserve -> insert (:2783-2796)
  void insert(std::vector<value_type> const &in) { this->insert_pairs(in.data(), in.data() + in.size(), true); }
  void insert(value_type const *b, value_type const *e) { this->insert_pairs(b, e, true); }
  void insert(std::vector<Key> const &in, T const &dflt) { this->insert_keys(in.data(), in.data() + in.size(), dflt, true); }
  void insert(Key const *b, Key const *e, T const &dflt) { this->insert_keys(b, e, dflt, true); }
  // insert_no_estimate(): grow by doubling only (:2802-2885)
  void insert_no_estimate(std::vector<value_type> const &in) { this->insert_pairs(in.data(), in.data() + in.size(), false); }
  void insert_no_estimate(value_type const *b, value_type const *e) { this->insert_pairs(b, e, false); }
  void insert_no_estimate(Key const *b, Key const *e, T const &dflt) { this->insert_keys(b, e, dflt, false); }
  void insert_no_estimate(std::vector<Key> const &in, T const &dflt) { this->insert_keys(in.data(), in.data() + in.size(), dflt, false); }

  // count: one 0/1 byte per query (:3311-3328)
  std::vector<uint8_t> count(Key const *b, Key const *e) const {
    std::vector<uint8_t> out(e - b);
    this->lookup_flags(b, e, out.data());
    return out;
  }
  size_t count(uint8_t *out, Key const *b, Key const *e) const { return this->lookup_flags(b, e, out); }
  bool exists(Key const &k) const { uint8_t r = 0; this->lookup_flags(&k, &k + 1, &r); return r != 0; }

  // find<OT>: OT = mapped_type or value_type, `nonexistent` fills the misses (:3376-3417)
  template <typename OT = T>
  typename std::enable_if<std::is_same<OT, T>::value, std::vector<OT>>::type
  find(Key const *b, Key const *e, T const &nonexistent = T()) const {
    std::vector<uint32_t> v(e - b);
    this->lookup_vals(b, e, v.data(), (uint32_t)nonexistent);
    return narrow(v);
  }
  template <typename OT>
  typename std::enable_if<std::is_same<OT, value_type>::value, std::vector<OT>>::type
  find(Key const *b, Key const *e, T const &nonexistent = T()) const {
    std::vector<uint32_t> v(e - b);
    this->lookup_vals(b, e, v.data(), (uint32_t)nonexistent);
    std::vector<value_type> out(e - b);
    for (size_t i = 0; i < out.size(); ++i) out[i] = value_type(b[i], static_cast<T>(v[i]));
    return out;
  }
  size_t find(T *out, Key const *b, Key const *e, T const &nonexistent = T()) const {
    std::vector<uint32_t> v(e - b);
    const size_t hits = this->lookup_vals(b, e, v.data(), (uint32_t)nonexistent);
    for (size_t i = 0; i < v.size(); ++i) out[i] = static_cast<T>(v[i]);
    return hits;
  }
  std::vector<value_type> find_existing(Key const *b, Key const *e) const { return this->hits(b, e); }

  size_t update(value_type const *b, value_type const *e) { return this->update_pairs(b, e); }
  size_t erase(Key const *b, Key const *e) { return this->erase_keys(b, e); }
  size_t erase(Key const &k) { return this->erase_keys(&k, &k + 1); }

 private:
  static std::vector<T> narrow(std::vector<uint32_t> const &v) {
    std::vector<T> out(v.size());
    for (size_t i = 0; i < v.size(); ++i) out[i] = static_cast<T>(v[i]);   // uint16_t counters: mod 2^16
    return out;
  }
};

template <typename Key, typename T, typename HashTag = hash::murmur3avx64, typename Equal = std::equal_to<Key>,
          typename Allocator = std::allocator<std::pair<const Key, T>>>
using hashmap_robinhood_offsets = hashmap_robinhood_offsets_reduction<Key, T, HashTag, Equal, DiscardReducer, Allocator>;   // :4094
template <typename Key, typename T, typename HashTag = hash::murmur3avx64, typename Equal = std::equal_to<Key>,
          typename Allocator = std::allocator<std::pair<const Key, T>>>
using hashmap_robinhood_offsets_count = hashmap_robinhood_offsets_reduction<Key, T, HashTag, Equal, std::plus<T>, Allocator>;   // :4099

// =====================================================================================================
// classic Robin Hood and linear-probe maps: insert never overwrites, count() returns size_type per key,
// find() returns the hits only (hashmap_robinhood.hpp:522-624,1102-1268,1294-1440;
// hashmap_linearprobe.hpp:430-513,580-889,911-1051).
// =====================================================================================================
namespace detail {
template <typename Key, typename T, typename HashTag, int MAPKIND>
class doubling_map : public table_base<Key, T, HashTag, KH_REDUCE_FIRST, MAPKIND> {
  using base = table_base<Key, T, HashTag, KH_REDUCE_FIRST, MAPKIND>;

 public:
  using typename base::value_type;
  using typename base::size_type;
  using base::base;

  void insert(std::vector<value_type> const &in) { this->insert_pairs(in.data(), in.data() + in.size(), false); }
  template <typename Iter> void insert(Iter b, Iter e) {
    std::vector<value_type> tmp(b, e);
    this->insert_pairs(tmp.data(), tmp.data() + tmp.size(), false);
  }
  void insert(value_type const *b, value_type const *e) { this->insert_pairs(b, e, false); }
  // single-element form: (true when the key was new) -- the iterator half of the reference's pair has no GPU meaning
  bool insert(value_type const &v) { return this->insert_pairs(&v, &v + 1, false) == 1; }

  template <typename Iter> std::vector<size_type> count(Iter b, Iter e) const {
    std::vector<Key> q = keys_of(b, e);
    std::vector<uint8_t> f(q.size());
    this->lookup_flags(q.data(), q.data() + q.size(), f.data());
    return std::vector<size_type>(f.begin(), f.end());
  }
  size_type count(Key const &k) const { uint8_t r = 0; this->lookup_flags(&k, &k + 1, &r); return r; }
  template <typename Iter> std::vector<value_type> find(Iter b, Iter e) const {
    std::vector<Key> q = keys_of(b, e);
    return this->hits(q.data(), q.data() + q.size());
  }
  template <typename Iter> size_t update(Iter b, Iter e) {
    std::vector<value_type> tmp(b, e);
    return this->update_pairs(tmp.data(), tmp.data() + tmp.size());
  }
  template <typename Iter> size_type erase(Iter b, Iter e) {
    std::vector<Key> q = keys_of(b, e);
    return this->erase_keys(q.data(), q.data() + q.size());
  }
  size_type erase(Key const &k) { return this->erase_keys(&k, &k + 1); }

 private:
  static Key const &key_of(Key const &k) { return k; }
  static Key const &key_of(value_type const &v) { return v.first; }
  template <typename Iter> static std::vector<Key> keys_of(Iter b, Iter e) {
    std::vector<Key> q;
    for (; b != e; ++b) q.push_back(key_of(*b));
    return q;
  }
};
}  // namespace detail

template <typename Key, typename T, typename HashTag = hash::murmur3avx64, typename Equal = std::equal_to<Key>,
          typename Allocator = std::allocator<std::pair<const Key, T>>>
using hashmap_robinhood_doubling = detail::doubling_map<Key, T, HashTag, KH_MAP_ROBINHOOD>;
template <typename Key, typename T, typename HashTag = hash::murmur3avx64, typename Equal = std::equal_to<Key>,
          typename Allocator = std::allocator<std::pair<const Key, T>>>
using hashmap_linearprobe_doubling = detail::doubling_map<Key, T, HashTag, KH_MAP_LINEARPROBE>;

// ---- free functions over key batches (parity surfaces; KH/hash_new.hpp batch mode, kmer_transform.hpp) ----
template <typename HashTag, typename Key>
inline std::vector<uint64_t> hash_batch(Key const *b, Key const *e, int device = 0) {
  std::vector<uint64_t> out(e - b);
  detail::ok(kh_hash_host(device, HashTag::kind, HashTag::seed, sizeof(Key) / 8, reinterpret_cast<const uint64_t *>(b), out.size(), out.data()), "hash");
  return out;
}
template <unsigned K>
inline void canonicalize(std::vector<Kmer<K>> &kmers, int device = 0) {   // lex_less in place
  uint64_t *p = reinterpret_cast<uint64_t *>(kmers.data());
  detail::ok(kh_canonicalize_host(device, K, p, kmers.size(), p), "canonicalize");
}
// KmerParser over N-split sequences: all (canonical) k-mers of the reads, in read order
template <unsigned K>
inline std::vector<Kmer<K>> kmers_from_reads(const char *bases, std::vector<uint64_t> const &offsets, bool canonical, int device = 0) {
  const uint64_t n_reads = offsets.empty() ? 0 : offsets.size() - 1;
  const int64_t n = detail::ok(kh_kmers_from_reads_host(device, bases, offsets.data(), n_reads, K, canonical, nullptr, 0), "kmers_from_reads");
  std::vector<Kmer<K>> out((size_t)n);
  if (n) detail::ok(kh_kmers_from_reads_host(device, bases, offsets.data(), n_reads, K, canonical, reinterpret_cast<uint64_t *>(out.data()), (uint64_t)n), "kmers_from_reads");
  return out;
}

}  // namespace khb
#endif  // KMERHASH_B200_HPP_
