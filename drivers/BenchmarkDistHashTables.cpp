// BenchmarkDistHashTables -- the reference's distributed hash table driver (benchmark/BenchmarkDistHashTables.cpp) over the
// single-process multi-GPU C ABI (khd_*): `-G n` GPUs of the box stand where `mpirun -np n` ranks stand for the reference.
//
// Same flags (:782-802): -C global element count, -R repeat rate, --missing-frac, --max_load / --min_load,
// --insert_prefetch / --query_prefetch (accepted, no meaning on a GPU); added: -G gpus, -I count|first|last for the three
// index flavours the reference builds as separate binaries (benchmarkHT-COUNT / -FIRST / -LAST, :976-1020), -S seed (the
// reference seeds from the clock, :389,502; a fixed seed makes runs comparable), --check to verify the results on the host.
// Same data recipe: unique 64-bit keys from mt19937 (:382-405), every key of the first half repeated floor(U(0, 2R)) + 1
// times, values 1, 2, 3, ... (:488-530), queries = input keys with a fraction of keys from the unused second half (:548-600).
// Same phases and report: init, copy_insert, insert, count, find, erase (benchmark(), :624-729).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "kmerhash_b200.hpp"

namespace {

struct Options {
  size_t count = 100000000;
  float repeats = 8.0f, missing_frac = 0.0f;
  double max_load = 0.8, min_load = 0.35;
  int gpus = 1, index = KH_REDUCE_COUNT;
  unsigned long seed = 1000;
  bool check = false;
};

struct PhaseTimer {
  std::chrono::steady_clock::time_point t0;
  std::vector<std::string> names; std::vector<double> secs; std::vector<size_t> counts;
  void start() { t0 = std::chrono::steady_clock::now(); }
  void end(const char *name, size_t n) {
    secs.push_back(std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
    names.push_back(name); counts.push_back(n);
  }
  void report(const char *title) const {
    std::printf("[TIME] %s\theader (s)\t[", title);
    for (auto &n : names) std::printf("%s,", n.c_str());
    std::printf("]\n[TIME] %s\tdur\t[", title);
    for (double s : secs) std::printf("%.6f,", s);
    std::printf("]\n[TIME] %s\tcnt\t[", title);
    for (size_t c : counts) std::printf("%zu,", c);
    std::printf("]\n[TIME] %s\tMops/s\t[", title);
    for (size_t i = 0; i < secs.size(); ++i) std::printf("%.1f,", secs[i] > 0 ? counts[i] / secs[i] / 1e6 : 0.0);
    std::printf("]\n");
  }
};

[[noreturn]] void usage(const char *argv0) {
  std::fprintf(stderr, "usage: %s [-C count] [-R repeat-rate] [--missing-frac f] [--max_load x] [--min_load x] [-G gpus] "
                       "[-I count|first|last] [-S seed] [--check]\n", argv0);
  std::exit(2);
}

Options parse(int argc, char **argv) {
  Options o;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto val = [&]() -> const char * { if (i + 1 >= argc) usage(argv[0]); return argv[++i]; };
    if (a == "-C" || a == "--count") o.count = std::strtoull(val(), nullptr, 10);
    else if (a == "-R" || a == "--repeat-rate") o.repeats = std::strtof(val(), nullptr);
    else if (a == "--missing-frac") o.missing_frac = std::strtof(val(), nullptr);
    else if (a == "--max_load") o.max_load = std::strtod(val(), nullptr);
    else if (a == "--min_load") o.min_load = std::strtod(val(), nullptr);
    else if (a == "--insert_prefetch" || a == "--query_prefetch") val();
    else if (a == "-G" || a == "--gpus") o.gpus = std::atoi(val());
    else if (a == "-S" || a == "--seed") o.seed = std::strtoul(val(), nullptr, 10);
    else if (a == "--check") o.check = true;
    else if (a == "-I" || a == "--index") {
      const std::string v = val();
      o.index = v == "first" ? KH_REDUCE_FIRST : v == "last" ? KH_REDUCE_LAST : v == "count" ? KH_REDUCE_COUNT : -1;
      if (o.index < 0) usage(argv[0]);
    } else usage(argv[0]);
  }
  if (o.repeats < 1.0f) o.repeats = 1.0f;
  return o;
}

}  // namespace

int main(int argc, char **argv) {
  const Options o = parse(argc, argv);
  using KeyType = uint64_t;
  using ValType = uint32_t;
  PhaseTimer data, test;

  // ---- data: get_unique_keys, generate_key_val_pairs, get_query (:382-600) ----
  data.start();
  std::unordered_set<KeyType> uniq;
  {
    std::mt19937 gen(o.seed);
    gen.discard(10000);
    std::uniform_int_distribution<KeyType> dist;
    const size_t want = std::max<size_t>(2, (size_t)((double)(o.count << 1) / o.repeats) + 2);   // two halves: inserted / missing
    uniq.reserve(want);
    while (uniq.size() < want) uniq.emplace(dist(gen));
  }
  data.end("get_keys", uniq.size());
  data.start();
  std::vector<KeyType> keys; std::vector<ValType> vals;
  std::vector<KeyType> missing;
  {
    std::mt19937 gen(o.seed + 1);
    gen.discard(10000);
    std::uniform_real_distribution<float> dist(0.0f, o.repeats * 2.0f);
    keys.reserve(o.count + (size_t)(o.repeats * 2.0f) + 1); vals.reserve(keys.capacity());
    auto it = uniq.begin();
    const size_t half = uniq.size() >> 1;
    size_t used = 0;
    ValType val = 1;
    while (keys.size() < o.count && used < half) {
      const size_t reps = o.repeats == 1.0f ? 1 : (size_t)dist(gen) + 1;
      for (size_t j = 0; j < reps; ++j, ++val) { keys.push_back(*it); vals.push_back(val); }
      ++it; ++used;
    }
    for (; it != uniq.end() && missing.size() < (size_t)(o.count * o.missing_frac) + 1; ++it) missing.push_back(*it);
    // std::random_shuffle of the pairs (:529): shuffle an index, apply it to both columns
    std::vector<uint32_t> idx(keys.size());
    for (size_t i = 0; i < idx.size(); ++i) idx[i] = (uint32_t)i;
    std::shuffle(idx.begin(), idx.end(), gen);
    std::vector<KeyType> k2(keys.size()); std::vector<ValType> v2(keys.size());
    for (size_t i = 0; i < idx.size(); ++i) { k2[i] = keys[idx[i]]; v2[i] = vals[idx[i]]; }
    keys.swap(k2); vals.swap(v2);
  }
  data.end("get_keyvals", keys.size());
  data.start();
  std::vector<KeyType> query(keys);
  {
    std::mt19937 gen(o.seed + 2);
    const size_t nmiss = (size_t)(query.size() * o.missing_frac);
    for (size_t i = 0; i < nmiss && !missing.empty(); ++i) query[gen() % query.size()] = missing[i % missing.size()];
  }
  data.end("get_query", query.size());
  data.report("data");

  // ---- benchmark() (:624-729) ----
  test.start();
  khd_map *map = khd_create(o.gpus, 1, KH_HASH_MURMUR64AVX, KH_HASH_MURMUR64AVX, o.index, 0);
  if (!map) { std::fprintf(stderr, "khd_create: %s\n", kh_last_error()); return 1; }
  khd_set_load_factors(map, o.min_load, o.max_load);
  test.end("init", khd_size(map));

  test.start();
  std::vector<KeyType> in(keys);                       // the reference inserts a copy (insert() permutes its argument)
  test.end("copy_insert", in.size());

  test.start();
  const int64_t added = o.index == KH_REDUCE_COUNT ? khd_insert_host(map, in.data(), in.size(), 1)
                                                   : khd_insert_pairs_host(map, in.data(), vals.data(), in.size());
  if (added < 0) { std::fprintf(stderr, "insert: %s\n", kh_last_error()); return 1; }
  test.end("insert", khd_size(map));
  std::printf("total map size after insert/rehash is %lu\n", (unsigned long)khd_size(map));

  std::vector<uint8_t> cres(query.size());
  test.start();
  const int64_t counted = khd_count_host(map, query.data(), query.size(), cres.data());
  test.end("count", (size_t)std::max<int64_t>(counted, 0));

  std::vector<ValType> fres(query.size());
  test.start();
  const int64_t found = khd_find_host(map, query.data(), query.size(), fres.data(), 0);
  test.end("find", (size_t)std::max<int64_t>(found, 0));

  int rc = 0;
  if (o.check) {                                        // host gold map: the reducer's semantics over the input order
    std::unordered_map<KeyType, ValType> gold;
    gold.reserve(uniq.size());
    for (size_t i = 0; i < keys.size(); ++i) {
      auto r = gold.emplace(keys[i], o.index == KH_REDUCE_COUNT ? 1u : vals[i]);
      if (!r.second) {
        if (o.index == KH_REDUCE_COUNT) ++r.first->second;
        else if (o.index == KH_REDUCE_LAST) r.first->second = vals[i];
      }
    }
    size_t bad = gold.size() == khd_size(map) ? 0 : 1;
    // FIRST / LAST on more than one GPU: the winner is decided by the exchange order (sender slice, then position), which for
    // slices of ONE input array in order is the input order -- the same rule the gold map applies
    for (size_t i = 0; i < query.size(); ++i) {
      auto it = gold.find(query[i]);
      const ValType exp = it == gold.end() ? 0u : it->second;
      if (fres[i] != exp || cres[i] != (it == gold.end() ? 0 : 1)) ++bad;
    }
    std::printf("check: %zu distinct keys, %zu mismatches\n", gold.size(), bad);
    rc = bad ? 1 : 0;
  }

  test.start();
  const int64_t erased = khd_erase_host(map, query.data(), query.size());
  if (erased < 0) { std::fprintf(stderr, "erase: %s\n", kh_last_error()); return 1; }
  test.end("erase", khd_size(map));
  test.report(o.index == KH_REDUCE_COUNT ? "benchmark:count" : o.index == KH_REDUCE_FIRST ? "benchmark:first" : "benchmark:last");
  if (o.check && o.missing_frac == 0.0f && khd_size(map) != 0) { std::printf("check: %lu keys left after erasing every key\n", (unsigned long)khd_size(map)); rc = 1; }
  khd_destroy(map);
  return rc;
}
