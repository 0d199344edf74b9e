// BenchmarkHashTables -- the reference's single-process table benchmark, driven through the
// khb:: facade (include/kmerhash_b200.hpp) instead of the CPU containers.
//
// Mirrors benchmark/BenchmarkHashTables.cpp of the reference: same options (-N -Q -R -m -c -k
// --max_load --min_load --insert_prefetch --query_prefetch -F), same input recipe (generate_input,
// :181-227: srand(23), two rand() per 64-bit word, sanitize, rand()%R extra copies,
// std::random_shuffle) and the same phase sequence and phase names (:1191-1364):
//   reserve, generate query, estimate, insert, find_pair, find_value, find_existing, count,
//   erase, count2
// Extra option: --count  -> hashmap_robinhood_offsets_count with insert(keys, 1) (the graded COUNT
// semantics) instead of the literal default (DiscardReducer, value = input index).
// Every phase is timed with std::chrono around the facade call: HOST vectors in, HOST vectors out.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "kmerhash_b200.hpp"
#include "generate_input.hpp"

namespace {

struct Options {
  size_t count = 100000000, query_frac = 2, repeats = 10;
  std::string map = "robinhood_offset_overflow", file, dump;
  bool canonical = false, count_mode = false;
  int k = 31;
  double max_load = 0.8, min_load = 0.35;
  unsigned insert_prefetch = 8, query_prefetch = 8;
};

Options parse(int argc, char **argv) {
  Options o;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto next = [&]() -> const char * {
      if (i + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", a.c_str()); std::exit(2); }
      return argv[++i];
    };
    if (a == "-N" || a == "--num_elements") o.count = std::strtoull(next(), nullptr, 10);
    else if (a == "-Q" || a == "--query_fraction") o.query_frac = std::strtoull(next(), nullptr, 10);
    else if (a == "-R" || a == "--repeate_rate") o.repeats = std::strtoull(next(), nullptr, 10);
    else if (a == "-m" || a == "--map_type") o.map = next();
    else if (a == "-F" || a == "--file") o.file = next();
    else if (a == "-c" || a == "--canonical") o.canonical = true;
    else if (a == "-k") o.k = std::atoi(next());
    else if (a == "--count") o.count_mode = true;
    else if (a == "--dump") o.dump = next();   // write the table after the insert phase in the -F format
    else if (a == "--max_load") o.max_load = std::atof(next());
    else if (a == "--min_load") o.min_load = std::atof(next());
    else if (a == "--insert_prefetch") o.insert_prefetch = std::atoi(next());
    else if (a == "--query_prefetch") o.query_prefetch = std::atoi(next());
    else if (a == "-h" || a == "--help") {
      std::puts("BenchmarkHashTables [-N n] [-Q frac] [-R repeats] [-k 31|63] [-c] [--count] [-F file]\n"
                "  [-m linearprobe|classic_robinhood|robinhood_offset|robinhood_offset_overflow]\n"
                "  [--max_load f] [--min_load f] [--insert_prefetch n] [--query_prefetch n]");
      std::exit(0);
    } else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); std::exit(2); }
  }
  if (o.query_frac == 0) o.query_frac = 1;
  return o;
}

struct PhaseTimer {
  std::chrono::steady_clock::time_point t0;
  std::vector<std::string> names;
  std::vector<double> secs;
  std::vector<size_t> counts;
  void start() { t0 = std::chrono::steady_clock::now(); }
  void end(const char *name, size_t n) {
    const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    names.push_back(name); secs.push_back(s); counts.push_back(n);
  }
  void report(const std::string &title) const {
    std::printf("[TIME] %s\theader (s)\t[", title.c_str());
    for (auto &n : names) std::printf("%s,", n.c_str());
    std::printf("]\n[TIME] %s\tdur\t[", title.c_str());
    for (double s : secs) std::printf("%.6f,", s);
    std::printf("]\n[TIME] %s\tcount\t[", title.c_str());
    for (size_t c : counts) std::printf("%zu,", c);
    std::printf("]\n[TIME] %s\tMops/s\t[", title.c_str());
    for (size_t i = 0; i < secs.size(); ++i) std::printf("%.1f,", secs[i] > 0 ? counts[i] / secs[i] / 1e6 : 0.0);
    std::printf("]\n");
  }
};

using khdrv::generate_input;   // generate_input.hpp (benchmark/BenchmarkHashTables.cpp:181-227)

// the -F format (include/kmerhash/io_utils.hpp:53-70): [u64 record bytes][u64 n][n records]
template <typename KmerT, typename Value>
bool load_file(const std::string &path, std::vector<std::pair<KmerT, Value>> &out) {
  FILE *f = std::fopen(path.c_str(), "rb");
  if (!f) return false;
  uint64_t rec = 0, n = 0;
  bool ok = std::fread(&rec, 8, 1, f) == 1 && std::fread(&n, 8, 1, f) == 1 && rec == sizeof(std::pair<KmerT, Value>);
  if (ok) { out.resize(n); ok = std::fread(out.data(), rec, n, f) == n; }
  std::fclose(f);
  return ok;
}

template <typename Pair>
void dump_file(const std::string &path, std::vector<Pair> const &v) {
  if (path.empty()) return;
  FILE *f = std::fopen(path.c_str(), "wb");
  if (!f) { std::fprintf(stderr, "cannot write %s\n", path.c_str()); std::exit(1); }
  const uint64_t rec = sizeof(Pair), n = v.size();
  std::fwrite(&rec, 8, 1, f); std::fwrite(&n, 8, 1, f);
  if (n) std::fwrite(v.data(), rec, n, f);
  std::fclose(f);
}

template <typename KmerT, typename Value>
std::vector<std::pair<KmerT, Value>> make_input(const Options &o, PhaseTimer &tm) {
  std::vector<std::pair<KmerT, Value>> input;
  tm.start();
  if (!o.file.empty()) {
    if (!load_file(o.file, input)) { std::fprintf(stderr, "cannot read %s\n", o.file.c_str()); std::exit(1); }
  } else {
    // -c: the reference canonicalises each unique key before copying it; canonicalising the whole
    // batch afterwards (one GPU call) yields the same vector
    generate_input(input, o.count, o.repeats);
    if (o.canonical) {
      std::vector<KmerT> ks(input.size());
      for (size_t i = 0; i < ks.size(); ++i) ks[i] = input[i].first;
      khb::canonicalize(ks);
      for (size_t i = 0; i < ks.size(); ++i) input[i].first = ks[i];
    }
  }
  tm.end("generate input", input.size());
  return input;
}

// benchmark_hashmap for the offsets family (:1191-1364)
template <typename MAP, typename KmerT, typename Value>
void benchmark_offsets(const std::string &name, const Options &o, std::vector<std::pair<KmerT, Value>> const &input,
                       PhaseTimer &tm, bool count_mode) {
  using value_type = typename MAP::value_type;
  tm.start();
  MAP map;
  map.set_max_load_factor(o.max_load);
  map.set_min_load_factor(o.min_load);
  map.set_insert_lookahead(o.insert_prefetch);
  map.set_query_lookahead(o.query_prefetch);
  tm.end("reserve", input.size());

  tm.start();
  std::vector<KmerT> query(input.size() / o.query_frac);
  for (size_t i = 0; i < query.size(); ++i) query[i] = input[i].first;
  tm.end("generate query", input.size());

  tm.start();
  std::vector<KmerT> keys(input.size());
  for (size_t i = 0; i < keys.size(); ++i) keys[i] = input[i].first;
  const double est = map.estimate(keys.data(), keys.size());
  tm.end("estimate", static_cast<size_t>(est));
  std::cout << "insert testing: estimated distinct = " << est << " in " << input.size() << std::endl;
  map.reserve(static_cast<size_t>(est));

  tm.start();
  if (count_mode) map.insert_no_estimate(keys.data(), keys.data() + keys.size(), Value(1));
  else map.insert_no_estimate(input.data(), input.data() + input.size());
  tm.end("insert", map.size());
  dump_file(o.dump, map.to_vector());

  tm.start();
  { auto finds = map.template find<value_type>(query.data(), query.data() + query.size()); tm.end("find_pair", finds.size()); }
  tm.start();
  { auto finds = map.template find<Value>(query.data(), query.data() + query.size()); tm.end("find_value", finds.size()); }
  tm.start();
  { auto finds = map.find_existing(query.data(), query.data() + query.size()); tm.end("find_existing", finds.size()); }
  tm.start();
  auto counts = map.count(query.data(), query.data() + query.size());
  tm.end("count", counts.size());
  size_t present = 0;
  for (auto c : counts) present += c;
  tm.start();
  const size_t erased = map.erase(query.data(), query.data() + query.size());
  tm.end("erase", erased);
  tm.start();
  counts = map.count(query.data(), query.data() + query.size());
  tm.end("count2", counts.size());
  size_t present2 = 0;
  for (auto c : counts) present2 += c;
  std::cout << name << ": size after erase " << map.size() << ", capacity " << map.capacity() << ", present "
            << present << ", erased " << erased << ", present after erase " << present2 << std::endl;
  tm.report(name);
}

// benchmark_hashmap_insert_mode for RH / LP (:588-780): per-map defaults, insert(vector), find hits, count as size_t
template <typename MAP, typename KmerT, typename Value>
void benchmark_doubling(const std::string &name, const Options &o, std::vector<std::pair<KmerT, Value>> const &input,
                        PhaseTimer &tm) {
  tm.start();
  MAP map;
  map.set_max_load_factor(o.max_load);
  map.set_min_load_factor(o.min_load);
  tm.end("reserve", input.size());
  tm.start();
  std::vector<KmerT> query(input.size() / o.query_frac);
  for (size_t i = 0; i < query.size(); ++i) query[i] = input[i].first;
  tm.end("generate query", input.size());
  tm.start();
  map.insert(input);
  tm.end("insert", map.size());
  dump_file(o.dump, map.to_vector());
  tm.start();
  { auto finds = map.find(query.begin(), query.end()); tm.end("find", finds.size()); }
  tm.start();
  auto counts = map.count(query.begin(), query.end());
  tm.end("count", counts.size());
  tm.start();
  const size_t erased = map.erase(query.begin(), query.end());
  tm.end("erase", erased);
  tm.start();
  counts = map.count(query.begin(), query.end());
  tm.end("count2", counts.size());
  std::cout << name << ": size after erase " << map.size() << ", capacity " << map.capacity() << ", erased " << erased << std::endl;
  tm.report(name);
}

template <unsigned K>
int run(const Options &o) {
  using KmerT = khb::Kmer<K>;
  using Value = uint32_t;
  using Hash = khb::hash::murmur3avx64;
  PhaseTimer tm;
  auto input = make_input<KmerT, Value>(o, tm);
  std::cout << " tuple size " << sizeof(std::pair<KmerT, Value>) << ", " << input.size() << " elements" << std::endl;
  if (o.map == "linearprobe")
    benchmark_doubling<khb::hashmap_linearprobe_doubling<KmerT, Value, Hash>>("hashmap_linearprobe_doubling", o, input, tm);
  else if (o.map == "classic_robinhood" || o.map == "robinhood")
    benchmark_doubling<khb::hashmap_robinhood_doubling<KmerT, Value, Hash>>("hashmap_robinhood_doubling", o, input, tm);
  else if (o.map == "robinhood_offset" || o.map == "robinhood_offset_overflow") {
    if (o.count_mode)
      benchmark_offsets<khb::hashmap_robinhood_offsets_count<KmerT, Value, Hash>>("hashmap_robinhood_offsets_count", o, input, tm, true);
    else
      benchmark_offsets<khb::hashmap_robinhood_offsets<KmerT, Value, Hash>>("hashmap_robinhood_offsets", o, input, tm, false);
  } else {
    std::fprintf(stderr, "map type '%s' is outside the GPU path (see DESIGN.md section 7)\n", o.map.c_str());
    return 2;
  }
  return 0;
}

}  // namespace

int main(int argc, char **argv) {
  const Options o = parse(argc, argv);
  try {
    if (o.k == 31) return run<31>(o);
    if (o.k == 63) return run<63>(o);
    if (o.k == 21) return run<21>(o);
    if (o.k == 32) return run<32>(o);
    std::fprintf(stderr, "k must be 21, 31, 32 or 63 in this driver\n");
    return 2;
  } catch (const std::exception &e) {
    std::fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
}
