// generate_input.hpp -- the benchmark drivers' input recipe, shared by drivers/BenchmarkHashTables.cpp and the ctypes
// helper bench.py loads (drivers/geninput.cpp), so that every arm of every measurement sees the reference's input bytes.
//
// Restates generate_input of the reference (benchmark/BenchmarkHashTables.cpp:181-227): srand(seed) (the reference: 23);
// per unique key every word = (rand() << 32) ^ rand(), sanitised; optionally the canonical strand; value = running index;
// rand() % repeats extra copies; std::random_shuffle (libstdc++: rand()-driven) at the end.  glibc rand() and libstdc++'s
// shuffle are called, not re-implemented, so the bytes are the reference's by construction.
#ifndef KH_DRIVERS_GENERATE_INPUT_HPP_
#define KH_DRIVERS_GENERATE_INPUT_HPP_
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <utility>
#include <vector>
#include "kmerhash_b200.hpp"

namespace khdrv {

// reverse complement of a 2-bit packed k-mer (KI/common/kmer.hpp:1078-1087): reverse the 2-bit groups of the whole
// word array, complement, shift the k significant groups down
template <unsigned K>
inline khb::Kmer<K> reverse_complement(khb::Kmer<K> const &k) {
  constexpr unsigned NW = khb::Kmer<K>::nWords;
  auto rev2 = [](uint64_t x) {
    x = ((x >> 2) & 0x3333333333333333ULL) | ((x & 0x3333333333333333ULL) << 2);
    x = ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL) | ((x & 0x0F0F0F0F0F0F0F0FULL) << 4);
    return __builtin_bswap64(x);
  };
  uint64_t r[NW];
  for (unsigned i = 0; i < NW; ++i) r[i] = ~rev2(k.data[NW - 1 - i]);
  const unsigned shift = 64 * NW - 2 * K;
  khb::Kmer<K> out;
  for (unsigned i = 0; i < NW; ++i) {
    uint64_t v = shift ? (r[i] >> shift) : r[i];
    if (shift && i + 1 < NW) v |= r[i + 1] << (64 - shift);
    out.data[i] = v;
  }
  out.sanitize();
  return out;
}

template <typename KmerT, typename Value>
void generate_input(std::vector<std::pair<KmerT, Value>> &out, size_t count, size_t repeats = 10, bool canonical = false,
                    unsigned seed = 23) {
  out.reserve(count);
  std::srand(seed);
  for (size_t i = 0; i < count; ++i) {
    KmerT k;
    for (unsigned j = 0; j < KmerT::nWords; ++j)
      k.data[j] = static_cast<uint64_t>(static_cast<long>(std::rand()) << 32) ^ static_cast<long>(std::rand());
    k.sanitize();
    if (canonical) {
      const KmerT kr = reverse_complement(k);
      k = (k < kr) ? k : kr;
    }
    out.emplace_back(k, static_cast<Value>(i));
    const size_t freq = std::rand() % repeats;
    for (size_t j = 0; j < freq; ++j)
      if (i + 1 < count) { ++i; out.emplace_back(k, static_cast<Value>(i)); }
  }
#if defined(__GNUC__)
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Wdeprecated-declarations"
#endif
  std::random_shuffle(out.begin(), out.end());   // libstdc++: rand()-driven, as in the reference
#if defined(__GNUC__)
#pragma GCC diagnostic pop
#endif
}

}  // namespace khdrv
#endif
