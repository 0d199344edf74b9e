// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE ONLY (never on the product path).
//
// A thin extern "C" wrapper that #includes the UNMODIFIED reference headers where
// they lie under /root/reference and exposes the reference's own CPU
// implementation of the hot path to the parity tests and to bench.py's
// cpu_baseline / --impl reference legs.  Built by oracle/Makefile into
// oracle/_ref/libkmerhash_ref.so.  No reference source is copied here: every
// function below only *calls* reference code.
//
// Reference entry points used (paths relative to /root/reference):
//   include/kmerhash/hash_new.hpp:136-351          identity/murmur32/murmur/murmur_x86/clhash/farm/farm32
//   include/kmerhash/murmurhash3_64_avx.hpp:1553    murmur3avx64
//   include/kmerhash/murmurhash3_32_avx.hpp:1252    murmur3avx32
//   include/kmerhash/murmurhash3_32_sse.hpp:1463    murmur3sse32
//   include/kmerhash/murmurhash3finalizer_32_avx.hpp:525  murmur3finalizer_avx32
//   include/kmerhash/crc32c_sse.hpp:121             crc32c
//   include/kmerhash/robinhood_offset_hashmap_ptr.hpp:196,4094,4099  offsets maps
//   include/kmerhash/hashmap_robinhood.hpp:126      hashmap_robinhood_doubling
//   include/kmerhash/hashmap_linearprobe.hpp:98     hashmap_linearprobe_doubling
//   include/kmerhash/hyperloglog64.hpp:145          hyperloglog64
//   ext/kmerind/src/common/kmer.hpp:115             bliss::common::Kmer
//   ext/kmerind/src/common/kmer_transform.hpp:109   lex_less
//
// The input generator restates benchmark/BenchmarkHashTables.cpp:181-227
// (generate_input) because that translation unit cannot be compiled without MPI;
// it calls the reference's Kmer::sanitize / reverse_complement / operator<.
//
// Build flags are value-relevant: -DNDEBUG (farmhash
// DebugTweak), SSE4.2+AES (farm32 == low 32 bits of farm), LZCNT (math_utils.hpp:78).

#include <cstdint>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <utility>
#include <algorithm>
#include <thread>
#include <chrono>
#include <functional>

#include "common/kmer.hpp"
#include "common/alphabets.hpp"
#include "common/kmer_transform.hpp"
#include "kmerhash/hash_new.hpp"
#include "kmerhash/hyperloglog64.hpp"
#include "kmerhash/robinhood_offset_hashmap_ptr.hpp"
#include "kmerhash/hashmap_robinhood.hpp"
#include "kmerhash/hashmap_linearprobe.hpp"

namespace {

using K31 = ::bliss::common::Kmer<31, ::bliss::common::DNA, uint64_t>;
using K63 = ::bliss::common::Kmer<63, ::bliss::common::DNA, uint64_t>;
using K21 = ::bliss::common::Kmer<21, ::bliss::common::DNA, uint64_t>;
using K15 = ::bliss::common::Kmer<15, ::bliss::common::DNA, uint64_t>;
using K47 = ::bliss::common::Kmer<47, ::bliss::common::DNA, uint64_t>;
using K33 = ::bliss::common::Kmer<33, ::bliss::common::DNA, uint64_t>;
using K5  = ::bliss::common::Kmer<5,  ::bliss::common::DNA, uint64_t>;
// other alphabets (benchmark/CMakeLists.txt:175-196: testKmerIndex 31 x {DNA, DNA5, DNA16}); DNA5 is the 3-bit DNA6 table
using K31_5  = ::bliss::common::Kmer<31, ::bliss::common::DNA5, uint64_t>;
using K21_5  = ::bliss::common::Kmer<21, ::bliss::common::DNA5, uint64_t>;
using K42_5  = ::bliss::common::Kmer<42, ::bliss::common::DNA5, uint64_t>;
using K7_5   = ::bliss::common::Kmer<7,  ::bliss::common::DNA5, uint64_t>;
using K31_16 = ::bliss::common::Kmer<31, ::bliss::common::DNA16, uint64_t>;
using K15_16 = ::bliss::common::Kmer<15, ::bliss::common::DNA16, uint64_t>;
using K32_16 = ::bliss::common::Kmer<32, ::bliss::common::DNA16, uint64_t>;
using K17_16 = ::bliss::common::Kmer<17, ::bliss::common::DNA16, uint64_t>;

static_assert(sizeof(K31) == 8,  "K31 must be one 64-bit word");
static_assert(sizeof(K63) == 16, "K63 must be two 64-bit words");

// hash kind ids: shared with include/kmerhash_b200.h (kh_hash_kind)
enum {
  H_IDEN = 0, H_MURMUR64 = 1, H_MURMUR64AVX = 2, H_MURMUR32 = 3, H_MURMUR32AVX = 4,
  H_MURMUR32SSE = 5, H_FARM = 6, H_FARM32 = 7, H_CLHASH = 8, H_CRC32C = 9,
  H_MURMUR32FIN = 10, H_MURMUR_X86 = 11
};

template <typename T> struct equal_to_ {
  inline constexpr bool operator()(T const & a, T const & b) const { return a == b; }
};

// ---- hashing -------------------------------------------------------------
template <typename H, typename KM>
void scalar_hash(H const & h, KM const * keys, size_t n, uint64_t * out) {
  for (size_t i = 0; i < n; ++i) out[i] = static_cast<uint64_t>(h(keys[i]));
}
template <typename H, typename KM, typename R>
void batch_hash(H const & h, KM const * keys, size_t n, uint64_t * out) {
  std::vector<R> tmp(n + 64);
  h(keys, n, tmp.data());
  for (size_t i = 0; i < n; ++i) out[i] = static_cast<uint64_t>(tmp[i]);
}

template <typename KM>
int hash_dispatch(int kind, uint64_t seed, KM const * keys, size_t n, uint64_t * out, int batch) {
  switch (kind) {
    case H_IDEN:       scalar_hash(::fsc::hash::identity<KM>(seed), keys, n, out); return 0;
    case H_MURMUR64:   scalar_hash(::fsc::hash::murmur<KM>(seed), keys, n, out); return 0;
    case H_MURMUR_X86: scalar_hash(::fsc::hash::murmur_x86<KM>(seed), keys, n, out); return 0;
    case H_MURMUR32:   scalar_hash(::fsc::hash::murmur32<KM>(seed), keys, n, out); return 0;
    case H_FARM:       scalar_hash(::fsc::hash::farm<KM>(seed), keys, n, out); return 0;
    case H_FARM32:     scalar_hash(::fsc::hash::farm32<KM>(seed), keys, n, out); return 0;
    case H_CLHASH:     scalar_hash(::fsc::hash::clhash<KM>(seed), keys, n, out); return 0;
    case H_MURMUR64AVX:
      if (batch) batch_hash<::fsc::hash::murmur3avx64<KM>, KM, uint64_t>(::fsc::hash::murmur3avx64<KM>(seed), keys, n, out);
      else scalar_hash(::fsc::hash::murmur3avx64<KM>(seed), keys, n, out);
      return 0;
    case H_MURMUR32AVX:
      if (batch) batch_hash<::fsc::hash::murmur3avx32<KM>, KM, uint32_t>(::fsc::hash::murmur3avx32<KM>(seed), keys, n, out);
      else scalar_hash(::fsc::hash::murmur3avx32<KM>(seed), keys, n, out);
      return 0;
    case H_MURMUR32SSE:
      if (batch) batch_hash<::fsc::hash::murmur3sse32<KM>, KM, uint32_t>(::fsc::hash::murmur3sse32<KM>(seed), keys, n, out);
      else scalar_hash(::fsc::hash::murmur3sse32<KM>(seed), keys, n, out);
      return 0;
    // murmur3finalizer_avx32 static_asserts sizeof(T) <= 4 (murmurhash3finalizer_32_avx.hpp:527):
    // not instantiable for 8/16-byte k-mer keys, so it is not part of this path.
    case H_CRC32C:
      if (batch) batch_hash<::fsc::hash::crc32c<KM>, KM, uint32_t>(::fsc::hash::crc32c<KM>(seed), keys, n, out);
      else scalar_hash(::fsc::hash::crc32c<KM>(seed), keys, n, out);
      return 0;
    default: return -1;
  }
}

// ---- k-mers --------------------------------------------------------------
template <typename KM>
size_t windows_impl(const char * s, size_t len, uint64_t * out) {
  // plain sliding window over every byte through the DNA LUT (no N / EOL logic):
  // exactly Kmer::nextFromChar(FROM_ASCII[c]) as in Kmer(std::string) kmer.hpp:241-248
  KM km;
  size_t n = 0;
  for (size_t i = 0; i < len; ++i) {
    km.nextFromChar(KM::KmerAlphabet::FROM_ASCII[static_cast<unsigned char>(s[i])]);
    if (i + 1 >= KM::size) {
      std::memcpy(out + n * KM::nWords, km.getData(), sizeof(KM));
      ++n;
    }
  }
  return n;
}
template <typename KM>
void revcomp_impl(const uint64_t * in, size_t n, uint64_t * out) {
  KM const * k = reinterpret_cast<KM const *>(in);
  KM * o = reinterpret_cast<KM *>(out);
  for (size_t i = 0; i < n; ++i) o[i] = k[i].reverse_complement();
}
template <typename KM>
void canonical_impl(const uint64_t * in, size_t n, uint64_t * out) {
  KM const * k = reinterpret_cast<KM const *>(in);
  KM * o = reinterpret_cast<KM *>(out);
  ::bliss::kmer::transform::lex_less<KM> canon;
  for (size_t i = 0; i < n; ++i) o[i] = canon(k[i]);
}

#define KH_DISPATCH_K(k, CALL)                    \
  switch (k) {                                    \
    case 5:  { using KM = K5;  CALL; } break;     \
    case 15: { using KM = K15; CALL; } break;     \
    case 21: { using KM = K21; CALL; } break;     \
    case 31: { using KM = K31; CALL; } break;     \
    case 33: { using KM = K33; CALL; } break;     \
    case 47: { using KM = K47; CALL; } break;     \
    case 63: { using KM = K63; CALL; } break;     \
    default: return -1;                           \
  }

// ---- maps ----------------------------------------------------------------
struct IMap {
  virtual ~IMap() {}
  virtual void insert_keys(const uint64_t * keys, size_t n, uint32_t dflt, int estimate) = 0;
  virtual void insert_pairs(const uint64_t * keys, const uint32_t * vals, size_t n, int estimate) = 0;
  virtual size_t count(const uint64_t * keys, size_t n, uint8_t * out) = 0;
  virtual size_t find(const uint64_t * keys, size_t n, uint32_t * out, uint32_t nonexistent) = 0;
  virtual size_t find_existing(const uint64_t * keys, size_t n, uint64_t * okeys, uint32_t * ovals) = 0;
  virtual size_t erase(const uint64_t * keys, size_t n) = 0;
  virtual size_t to_vector(uint64_t * okeys, uint32_t * ovals, size_t cap) = 0;
  virtual size_t size() = 0;
  virtual size_t capacity() = 0;
  virtual void reserve(size_t n) = 0;
  virtual double hll_estimate(const uint64_t * keys, size_t n) = 0;
};

// offsets maps (batch API, robinhood_offset_hashmap_ptr.hpp:2783-2885,3251-3454,3781)
template <typename MAP, typename KM>
struct OffsetsMap : public IMap {
  MAP m;
  static constexpr size_t W = KM::nWords;
  OffsetsMap(double minl, double maxl) { m.set_min_load_factor(minl); m.set_max_load_factor(maxl); }
  void insert_keys(const uint64_t * keys, size_t n, uint32_t dflt, int estimate) override {
    KM const * k = reinterpret_cast<KM const *>(keys);
    if (estimate) m.insert(k, k + n, dflt); else m.insert_no_estimate(k, k + n, dflt);
  }
  void insert_pairs(const uint64_t * keys, const uint32_t * vals, size_t n, int estimate) override {
    std::vector<std::pair<KM, uint32_t>> in(n);
    KM const * k = reinterpret_cast<KM const *>(keys);
    for (size_t i = 0; i < n; ++i) { in[i].first = k[i]; in[i].second = vals[i]; }
    if (estimate) m.insert(in); else m.insert_no_estimate(in);
  }
  size_t count(const uint64_t * keys, size_t n, uint8_t * out) override {
    KM * k = const_cast<KM *>(reinterpret_cast<KM const *>(keys));
    return m.count(out, k, k + n);
  }
  size_t find(const uint64_t * keys, size_t n, uint32_t * out, uint32_t nonexistent) override {
    KM * k = const_cast<KM *>(reinterpret_cast<KM const *>(keys));
    return m.find(out, k, k + n, nonexistent);
  }
  size_t find_existing(const uint64_t * keys, size_t n, uint64_t * okeys, uint32_t * ovals) override {
    KM * k = const_cast<KM *>(reinterpret_cast<KM const *>(keys));
    auto res = m.find_existing(k, k + n);
    for (size_t i = 0; i < res.size(); ++i) {
      std::memcpy(okeys + i * W, res[i].first.getData(), sizeof(KM));
      ovals[i] = res[i].second;
    }
    return res.size();
  }
  size_t erase(const uint64_t * keys, size_t n) override {
    KM * k = const_cast<KM *>(reinterpret_cast<KM const *>(keys));
    return m.erase(k, k + n);
  }
  size_t to_vector(uint64_t * okeys, uint32_t * ovals, size_t cap) override {
    auto v = m.to_vector();
    size_t c = std::min(cap, v.size());
    for (size_t i = 0; i < c; ++i) {
      std::memcpy(okeys + i * W, v[i].first.getData(), sizeof(KM));
      ovals[i] = v[i].second;
    }
    return v.size();
  }
  size_t size() override { return m.size(); }
  size_t capacity() override { return m.capacity(); }
  void reserve(size_t n) override { m.reserve(n); }
  double hll_estimate(const uint64_t * keys, size_t n) override {
    // the driver's "estimate" phase, benchmark/BenchmarkHashTables.cpp:1232-1240
    KM const * k = reinterpret_cast<KM const *>(keys);
    m.get_hll().update(k, n);
    return m.get_hll().estimate();
  }
};

// classic RH / LP maps (per-key API, first insert wins; hashmap_robinhood.hpp:522-717,
// hashmap_linearprobe.hpp:430-573)
template <typename MAP, typename KM>
struct ClassicMap : public IMap {
  MAP m;
  static constexpr size_t W = KM::nWords;
  ClassicMap(double minl, double maxl) : m(128, minl, maxl) {}
  void insert_keys(const uint64_t * keys, size_t n, uint32_t dflt, int) override {
    KM const * k = reinterpret_cast<KM const *>(keys);
    for (size_t i = 0; i < n; ++i) m.insert(k[i], dflt);
  }
  void insert_pairs(const uint64_t * keys, const uint32_t * vals, size_t n, int) override {
    KM const * k = reinterpret_cast<KM const *>(keys);
    for (size_t i = 0; i < n; ++i) m.insert(k[i], vals[i]);
  }
  size_t count(const uint64_t * keys, size_t n, uint8_t * out) override {
    KM const * k = reinterpret_cast<KM const *>(keys);
    size_t c = 0;
    for (size_t i = 0; i < n; ++i) { out[i] = static_cast<uint8_t>(m.count(k[i])); c += out[i]; }
    return c;
  }
  size_t find(const uint64_t * keys, size_t n, uint32_t * out, uint32_t nonexistent) override {
    KM const * k = reinterpret_cast<KM const *>(keys);
    size_t c = 0;
    auto e = m.end();
    for (size_t i = 0; i < n; ++i) {
      auto it = m.find(k[i]);
      if (it != e) { out[i] = (*it).second; ++c; } else out[i] = nonexistent;
    }
    return c;
  }
  size_t find_existing(const uint64_t * keys, size_t n, uint64_t * okeys, uint32_t * ovals) override {
    KM const * k = reinterpret_cast<KM const *>(keys);
    size_t c = 0;
    auto e = m.end();
    for (size_t i = 0; i < n; ++i) {
      auto it = m.find(k[i]);
      if (it != e) {
        std::memcpy(okeys + c * W, (*it).first.getData(), sizeof(KM));
        ovals[c] = (*it).second; ++c;
      }
    }
    return c;
  }
  size_t erase(const uint64_t * keys, size_t n) override {
    KM const * k = reinterpret_cast<KM const *>(keys);
    return m.erase(k, k + n);
  }
  size_t to_vector(uint64_t * okeys, uint32_t * ovals, size_t cap) override {
    auto v = m.to_vector();
    size_t c = std::min(cap, v.size());
    for (size_t i = 0; i < c; ++i) {
      std::memcpy(okeys + i * W, v[i].first.getData(), sizeof(KM));
      ovals[i] = v[i].second;
    }
    return v.size();
  }
  size_t size() override { return m.size(); }
  size_t capacity() override { return cap_of(m, 0); }
  void reserve(size_t n) override { m.reserve(n); }
  double hll_estimate(const uint64_t *, size_t) override { return -1.0; }
 private:
  // hashmap_linearprobe_doubling exposes no capacity(); report 0 for it.
  template <typename M> static auto cap_of(M & mm, int) -> decltype(mm.capacity()) { return mm.capacity(); }
  template <typename M> static size_t cap_of(M &, long) { return 0; }
};

enum { M_OFFSETS_COUNT = 0, M_OFFSETS_FIRST = 1, M_OFFSETS_LAST = 2, M_ROBINHOOD = 3, M_LINEARPROBE = 4 };

template <typename KM, template <typename> class HASH>
IMap * make_map_h(int map_kind, double minl, double maxl) {
  switch (map_kind) {
    case M_OFFSETS_COUNT:
      return new OffsetsMap<::fsc::hashmap_robinhood_offsets_count<KM, uint32_t, HASH, equal_to_>, KM>(minl, maxl);
    case M_OFFSETS_FIRST:
      return new OffsetsMap<::fsc::hashmap_robinhood_offsets<KM, uint32_t, HASH, equal_to_>, KM>(minl, maxl);
    case M_OFFSETS_LAST:
      return new OffsetsMap<::fsc::hashmap_robinhood_offsets_reduction<KM, uint32_t, HASH, equal_to_, ::fsc::ReplaceReducer>, KM>(minl, maxl);
    default: return nullptr;
  }
}

template <typename KM>
IMap * make_map(int map_kind, int hash_kind, double minl, double maxl) {
  if (map_kind == M_ROBINHOOD)
    return new ClassicMap<::fsc::hashmap_robinhood_doubling<KM, uint32_t, ::fsc::hash::murmur<KM>, equal_to_<KM>>, KM>(minl, maxl);
  if (map_kind == M_LINEARPROBE)
    return new ClassicMap<::fsc::hashmap_linearprobe_doubling<KM, uint32_t, ::fsc::hash::murmur<KM>, equal_to_<KM>>, KM>(minl, maxl);
  switch (hash_kind) {
    case H_MURMUR64:    return make_map_h<KM, ::fsc::hash::murmur>(map_kind, minl, maxl);
    case H_MURMUR64AVX: return make_map_h<KM, ::fsc::hash::murmur3avx64>(map_kind, minl, maxl);
    case H_FARM:        return make_map_h<KM, ::fsc::hash::farm>(map_kind, minl, maxl);
#ifndef KH_REF_FEWER_MAPS
    case H_CLHASH:      return make_map_h<KM, ::fsc::hash::clhash>(map_kind, minl, maxl);
    case H_CRC32C:      return make_map_h<KM, ::fsc::hash::crc32c>(map_kind, minl, maxl);
    case H_MURMUR32AVX: return make_map_h<KM, ::fsc::hash::murmur3avx32>(map_kind, minl, maxl);
#endif
    default: return nullptr;
  }
}

// ---- rank assignment (distributed_batched_robinhood_map.hpp:164-178,652,749) ----
template <typename KM, typename H>
void rank_impl(H const & h, KM const * keys, size_t n, uint32_t p, uint32_t * out) {
  using HV = decltype(h(keys[0]));
  bool pow2 = (p & (p - 1)) == 0;
  HV cnt = static_cast<HV>(p - (pow2 ? 1 : 0));
  for (size_t i = 0; i < n; ++i) {
    HV x = h(keys[i]);
    out[i] = static_cast<uint32_t>(pow2 ? (x & cnt) : (x % cnt));
  }
}
template <typename KM>
int rank_dispatch(int hash_kind, KM const * keys, size_t n, uint32_t p, uint32_t * out) {
  const uint64_t seed = 9876543;
  switch (hash_kind) {
    case H_IDEN:        rank_impl(::fsc::hash::identity<KM>(seed), keys, n, p, out); return 0;
    case H_MURMUR64:    rank_impl(::fsc::hash::murmur<KM>(seed), keys, n, p, out); return 0;
    case H_MURMUR64AVX: rank_impl(::fsc::hash::murmur3avx64<KM>(seed), keys, n, p, out); return 0;
    case H_MURMUR32:    rank_impl(::fsc::hash::murmur32<KM>(seed), keys, n, p, out); return 0;
    case H_MURMUR32AVX: rank_impl(::fsc::hash::murmur3avx32<KM>(seed), keys, n, p, out); return 0;
    case H_FARM:        rank_impl(::fsc::hash::farm<KM>(seed), keys, n, p, out); return 0;
    case H_FARM32:      rank_impl(::fsc::hash::farm32<KM>(seed), keys, n, p, out); return 0;
    case H_CLHASH:      rank_impl(::fsc::hash::clhash<KM>(seed), keys, n, p, out); return 0;
    case H_CRC32C:      rank_impl(::fsc::hash::crc32c<KM>(seed), keys, n, p, out); return 0;
    default: return -1;
  }
}

// ---- input generator: restates benchmark/BenchmarkHashTables.cpp:181-227 ----
template <typename KM>
size_t generate_input_impl(size_t count, size_t repeats, int canonical, uint64_t * okeys, uint32_t * ovals) {
  std::vector<std::pair<KM, uint32_t>> output;
  output.reserve(count);
  KM k, kr;
  srand(23);
  for (size_t i = 0; i < count; ++i) {
    for (size_t j = 0; j < KM::nWords; ++j)
      k.getDataRef()[j] = static_cast<typename KM::KmerWordType>(static_cast<long>(rand()) << 32) ^ static_cast<long>(rand());
    k.sanitize();
    if (canonical) { kr = k.reverse_complement(); k = (k < kr) ? k : kr; }
    output.emplace_back(k, static_cast<uint32_t>(i));
    size_t freq = rand() % repeats;
    for (size_t j = 0; j < freq; ++j) {
      if (i + 1 < count) { ++i; output.emplace_back(k, static_cast<uint32_t>(i)); }
    }
  }
  std::random_shuffle(output.begin(), output.end());
  for (size_t i = 0; i < output.size(); ++i) {
    std::memcpy(okeys + i * KM::nWords, output[i].first.getData(), sizeof(KM));
    ovals[i] = output[i].second;
  }
  return output.size();
}

#define KH_DISPATCH_AK(alpha, k, CALL)                                  \
  switch ((alpha) * 1000 + (k)) {                                        \
    case 5031:  { using KM = K31_5;  CALL; } break;                      \
    case 5021:  { using KM = K21_5;  CALL; } break;                      \
    case 5042:  { using KM = K42_5;  CALL; } break;                      \
    case 5007:  { using KM = K7_5;   CALL; } break;                      \
    case 16031: { using KM = K31_16; CALL; } break;                      \
    case 16015: { using KM = K15_16; CALL; } break;                      \
    case 16032: { using KM = K32_16; CALL; } break;                      \
    case 16017: { using KM = K17_16; CALL; } break;                      \
    default: return -1;                                                  \
  }

}  // namespace

extern "C" {

// alphabet = 5 (DNA5 = the 3-bit DNA6 table, KI/common/alphabets.hpp:197-292,746) or 16 (DNA16, :648-739)
long ref_kmer_windows_alpha(int alpha, int k, const char * s, size_t len, uint64_t * out) {
  size_t n = 0;
  KH_DISPATCH_AK(alpha, k, n = windows_impl<KM>(s, len, out));
  return static_cast<long>(n);
}
int ref_revcomp_alpha(int alpha, int k, const uint64_t * in, size_t n, uint64_t * out) {
  KH_DISPATCH_AK(alpha, k, revcomp_impl<KM>(in, n, out));
  return 0;
}
int ref_canonical_alpha(int alpha, int k, const uint64_t * in, size_t n, uint64_t * out) {
  KH_DISPATCH_AK(alpha, k, canonical_impl<KM>(in, n, out));
  return 0;
}
int ref_kmer_words_alpha(int alpha, int k) {
  int w = -1;
  KH_DISPATCH_AK(alpha, k, w = KM::nWords);
  return w;
}

int ref_hash(int kind, uint64_t seed, int key_words, const uint64_t * keys, size_t n, uint64_t * out, int batch) {
  if (key_words == 1) return hash_dispatch(kind, seed, reinterpret_cast<K31 const *>(keys), n, out, batch);
  if (key_words == 2) return hash_dispatch(kind, seed, reinterpret_cast<K63 const *>(keys), n, out, batch);
  return -1;
}

long ref_kmer_windows(int k, const char * s, size_t len, uint64_t * out) {
  size_t n = 0;
  KH_DISPATCH_K(k, n = windows_impl<KM>(s, len, out));
  return static_cast<long>(n);
}
int ref_revcomp(int k, const uint64_t * in, size_t n, uint64_t * out) {
  KH_DISPATCH_K(k, revcomp_impl<KM>(in, n, out));
  return 0;
}
int ref_canonical(int k, const uint64_t * in, size_t n, uint64_t * out) {
  KH_DISPATCH_K(k, canonical_impl<KM>(in, n, out));
  return 0;
}
int ref_kmer_words(int k) {
  int w = -1;
  KH_DISPATCH_K(k, w = KM::nWords);
  return w;
}

void * ref_map_create(int map_kind, int key_words, int hash_kind, double min_load, double max_load) {
  if (key_words == 1) return make_map<K31>(map_kind, hash_kind, min_load, max_load);
  if (key_words == 2) return make_map<K63>(map_kind, hash_kind, min_load, max_load);
  return nullptr;
}
void ref_map_destroy(void * m) { delete static_cast<IMap *>(m); }
void ref_map_insert_keys(void * m, const uint64_t * keys, size_t n, uint32_t dflt, int estimate) {
  static_cast<IMap *>(m)->insert_keys(keys, n, dflt, estimate);
}
void ref_map_insert_pairs(void * m, const uint64_t * keys, const uint32_t * vals, size_t n, int estimate) {
  static_cast<IMap *>(m)->insert_pairs(keys, vals, n, estimate);
}
size_t ref_map_count(void * m, const uint64_t * keys, size_t n, uint8_t * out) { return static_cast<IMap *>(m)->count(keys, n, out); }
size_t ref_map_find(void * m, const uint64_t * keys, size_t n, uint32_t * out, uint32_t nonexistent) {
  return static_cast<IMap *>(m)->find(keys, n, out, nonexistent);
}
size_t ref_map_find_existing(void * m, const uint64_t * keys, size_t n, uint64_t * okeys, uint32_t * ovals) {
  return static_cast<IMap *>(m)->find_existing(keys, n, okeys, ovals);
}
size_t ref_map_erase(void * m, const uint64_t * keys, size_t n) { return static_cast<IMap *>(m)->erase(keys, n); }
size_t ref_map_to_vector(void * m, uint64_t * okeys, uint32_t * ovals, size_t cap) { return static_cast<IMap *>(m)->to_vector(okeys, ovals, cap); }
size_t ref_map_size(void * m) { return static_cast<IMap *>(m)->size(); }
size_t ref_map_capacity(void * m) { return static_cast<IMap *>(m)->capacity(); }
void ref_map_reserve(void * m, size_t n) { static_cast<IMap *>(m)->reserve(n); }
double ref_map_hll_estimate(void * m, const uint64_t * keys, size_t n) { return static_cast<IMap *>(m)->hll_estimate(keys, n); }

int ref_rank(int hash_kind, int key_words, const uint64_t * keys, size_t n, uint32_t p, uint32_t * out) {
  if (p == 0) return -1;
  if (key_words == 1) return rank_dispatch(hash_kind, reinterpret_cast<K31 const *>(keys), n, p, out);
  if (key_words == 2) return rank_dispatch(hash_kind, reinterpret_cast<K63 const *>(keys), n, p, out);
  return -1;
}

size_t ref_generate_input(int key_words, size_t count, size_t repeats, int canonical, uint64_t * okeys, uint32_t * ovals) {
  if (key_words == 1) return generate_input_impl<K31>(count, repeats, canonical, okeys, ovals);
  if (key_words == 2) return generate_input_impl<K63>(count, repeats, canonical, okeys, ovals);
  return 0;
}

// glibc rand() probe used to pin the clhash key table (hash_new.hpp:290-296)
void ref_rand_words(unsigned seed, int n, uint64_t * out) {
  srand(seed);
  for (int i = 0; i < n; ++i)
    out[i] = (static_cast<unsigned long>(rand()) << 32) | static_cast<unsigned long>(rand());
}

// ---- timed CPU baseline ---------------------------------------------------
// Replays benchmark_hashmap's phases (benchmark/BenchmarkHashTables.cpp:1191-1364)
// on `threads` independent shards.  With threads > 1 each shard owns the keys whose
// reference rank (DistHash seed 9876543, & (P-1) or % P) equals the shard id, i.e.
// the per-rank work of BenchmarkDistHashTables WITHOUT the MPI all-to-all (MPI is
// not available).  Partitioning is timed separately.
// out_secs: [0]=partition, [1]=estimate+reserve, [2]=insert, [3]=find, [4]=count, [5]=erase, [6]=count2
// out_counts: [0]=distinct after insert, [1]=found, [2]=counted, [3]=erased, [4]=counted2
int ref_baseline_run(int map_kind, int key_words, int hash_kind, double min_load, double max_load,
                     const uint64_t * keys, size_t n, const uint64_t * queries, size_t nq,
                     int threads, double * out_secs, uint64_t * out_counts) {
  using clk = std::chrono::steady_clock;
  auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
  if (threads < 1) threads = 1;
  const size_t W = static_cast<size_t>(key_words);
  for (int i = 0; i < 7; ++i) out_secs[i] = 0;
  for (int i = 0; i < 5; ++i) out_counts[i] = 0;

  std::vector<std::vector<uint64_t>> sk(threads), sq(threads);
  auto t0 = clk::now();
  if (threads == 1) {
    sk[0].assign(keys, keys + n * W);
    sq[0].assign(queries, queries + nq * W);
    t0 = clk::now();
  } else {
    // one counting pass + one scatter, both parallel over input ranges: thread t ranks its range, counts it per shard, and
    // (after a prefix over threads) copies its keys to their place inside every shard -- O(n) in total
    auto partition = [&](const uint64_t * src, size_t cnt, std::vector<std::vector<uint64_t>> & dst) {
      std::vector<uint32_t> r(cnt);
      std::vector<size_t> c((size_t)threads * threads, 0);          // c[t * threads + s]: keys of range t for shard s
      auto par = [&](std::function<void(int, size_t, size_t)> f) {
        std::vector<std::thread> th;
        for (int t = 0; t < threads; ++t) th.emplace_back([&, t]() { f(t, cnt * t / threads, cnt * (t + 1) / threads); });
        for (auto & x : th) x.join();
      };
      par([&](int t, size_t b, size_t e) {
        ref_rank(hash_kind, key_words, src + b * W, e - b, threads, r.data() + b);
        for (size_t i = b; i < e; ++i) ++c[(size_t)t * threads + r[i]];
      });
      for (int s = 0; s < threads; ++s) {
        size_t acc = 0;
        for (int t = 0; t < threads; ++t) { const size_t v = c[(size_t)t * threads + s]; c[(size_t)t * threads + s] = acc; acc += v; }
        dst[s].resize(acc * W);
      }
      par([&](int t, size_t b, size_t e) {
        size_t * at = &c[(size_t)t * threads];
        for (size_t i = b; i < e; ++i) {
          uint64_t * o = dst[r[i]].data() + at[r[i]]++ * W;
          for (size_t w = 0; w < W; ++w) o[w] = src[i * W + w];
        }
      });
    };
    partition(keys, n, sk);
    partition(queries, nq, sq);
  }
  auto t1 = clk::now();
  out_secs[0] = (threads == 1) ? 0.0 : secs(t0, t1);

  std::vector<IMap *> maps(threads, nullptr);
  for (int t = 0; t < threads; ++t) {
    maps[t] = static_cast<IMap *>(ref_map_create(map_kind, key_words, hash_kind, min_load, max_load));
    if (!maps[t]) return -1;
  }
  std::vector<uint64_t> cnt(threads * 5, 0);
  auto phase = [&](int idx, std::function<void(int)> f) {
    auto a = clk::now();
    std::vector<std::thread> th;
    for (int t = 0; t < threads; ++t) th.emplace_back(f, t);
    for (auto & x : th) x.join();
    out_secs[idx] = secs(a, clk::now());
  };
  phase(1, [&](int t) {
    size_t m = sk[t].size() / W;
    if (map_kind <= M_OFFSETS_LAST) { double est = maps[t]->hll_estimate(sk[t].data(), m); maps[t]->reserve(static_cast<size_t>(est)); }
  });
  phase(2, [&](int t) { maps[t]->insert_keys(sk[t].data(), sk[t].size() / W, 1, 0); cnt[t * 5 + 0] = maps[t]->size(); });
  phase(3, [&](int t) { size_t m = sq[t].size() / W; std::vector<uint32_t> o(m + 1); cnt[t * 5 + 1] = maps[t]->find(sq[t].data(), m, o.data(), 0); });
  phase(4, [&](int t) { size_t m = sq[t].size() / W; std::vector<uint8_t> o(m + 1); cnt[t * 5 + 2] = maps[t]->count(sq[t].data(), m, o.data()); });
  phase(5, [&](int t) { cnt[t * 5 + 3] = maps[t]->erase(sq[t].data(), sq[t].size() / W); });
  phase(6, [&](int t) { size_t m = sq[t].size() / W; std::vector<uint8_t> o(m + 1); cnt[t * 5 + 4] = maps[t]->count(sq[t].data(), m, o.data()); });
  for (int t = 0; t < threads; ++t) {
    for (int j = 0; j < 5; ++j) out_counts[j] += cnt[t * 5 + j];
    delete maps[t];
  }
  return 0;
}

}  // extern "C"
