// geninput.cpp -- libkh_geninput.so: the drivers' generate_input (generate_input.hpp) behind a C entry point, so bench.py
// and the tests feed every arm the bytes the reference's BenchmarkHashTables generates.  Host only, no CUDA.
#include <cstring>
#include "generate_input.hpp"

namespace {
template <unsigned K>
size_t run(size_t count, size_t repeats, int canonical, unsigned seed, uint64_t *keys, uint32_t *vals) {
  using KmerT = khb::Kmer<K>;
  std::vector<std::pair<KmerT, uint32_t>> v;
  khdrv::generate_input(v, count, repeats, canonical != 0, seed);
  for (size_t i = 0; i < v.size(); ++i) {
    std::memcpy(keys + i * KmerT::nWords, v[i].first.data, sizeof(uint64_t) * KmerT::nWords);
    if (vals) vals[i] = v[i].second;
  }
  return v.size();
}
}  // namespace

// k = 31 (8-byte keys) or 63 (16-byte keys); keys: count * nWords words, vals: count words or null.  Returns the number
// of records written (== count).  Not thread-safe: drives the process-wide rand().
extern "C" size_t khg_generate_input(int k, size_t count, size_t repeats, int canonical, unsigned seed, uint64_t *keys, uint32_t *vals) {
  if (!keys || repeats == 0) return 0;
  if (k == 31) return run<31>(count, repeats, canonical, seed, keys, vals);
  if (k == 63) return run<63>(count, repeats, canonical, seed, keys, vals);
  return 0;
}
