// BenchmarkKmerCounter -- the reference's k-mer counting driver (benchmark/BenchmarkKmerCounter.cpp) over the
// khb:: facade and the multi-GPU C ABI (khd_*), instead of the MPI count index.
//
// Same shape as the reference: positional FASTA / FASTQ file names (or --synthetic for the BASELINE.json
// config-4 generator), canonical k-mers (the shipped targets are ...-CANONICAL-...-COUNT-dtIDEN-dhMURMUR64avx-
// shMURMUR64avx, benchmark/CMakeLists.txt:325-332), CountType = uint16_t (:184), phases
//   init, read, parse (KmerParser = kh_kmers_from_reads), insert, [count, find of a sample], write
// and -O writes [Kmer][count] records like the reference's writers (:1802-2050): 8-byte k-mer word + 2-byte count.
// -G n runs the count map hash-partitioned over n GPUs of the box (owner = murmur3avx64_{9876543}(kmer) & (n-1)),
// which is what `mpirun -np n` is to the reference.
//
// File reading is partitioned like the reference's (byte ranges, record-boundary search, k - 1 overlap for FASTA sequences
// that cross a range border; host threads stand where MPI ranks stand); the k-mer generation kernel itself skips CR/LF and
// splits at N.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <random>
#include <string>
#include <thread>
#include <vector>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "kmerhash_b200.hpp"

namespace {

constexpr unsigned K = 31;
using KmerT = khb::Kmer<K>;
using CountType = uint16_t;

struct Options {
  std::vector<std::string> files;
  std::string out;
  size_t syn_reads = 0, syn_len = 150, syn_genome = 100000000;
  int gpus = 1;
  size_t sample = 1000000;
  std::string reads_only;      // test hook: write what the reader produced ([u64 n_reads][offsets][bases]) and stop
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
    std::printf("]\n");
  }
};

// ---- partitioned file reading -----------------------------------------------------------------------------------------------
// The reference hands every MPI rank a byte range of the file (KI/io/kmer_file_helper.hpp:407-645 read_file_mmap / _posix /
// _mpiio): a rank looks for the first record boundary inside its range (FASTQ: a line starting with '@' whose next-but-one
// line starts with '+', KI/io/fastq_loader.hpp; FASTA: a line starting with '>', KI/io/fasta_loader.hpp), parses the records
// that START in its range, and a FASTA sequence that runs across a range border is cut there with the next k - 1 bases
// repeated, so that every k-mer is produced exactly once (KI/io/filtered_sequence_iterator.hpp:179-440).  Same scheme here with
// host threads in the role of the ranks: the file is mapped, split into byte ranges, every thread fills its own
// (bases, offsets) piece, the pieces are concatenated in order -- the k-mers and their order are those of a serial parse.
// FASTA pieces keep their line breaks (the kernel filters EOL like the reference's iterator, KI/utils/file_utils.hpp:43-47).
struct Piece { std::string bases; std::vector<uint64_t> ends; };   // ends[i] = end of read i inside `bases`

inline const char *line_end(const char *p, const char *e) { const void *q = memchr(p, '\n', (size_t)(e - p)); return q ? (const char *)q : e; }
inline const char *next_line(const char *p, const char *e) { const char *q = line_end(p, e); return q < e ? q + 1 : e; }

// first FASTQ record start at or behind `from` (a line start)
const char *fastq_record_start(const char *base, const char *from, const char *e) {
  const char *p = from;
  if (p != base) p = next_line(p - 1, e);                 // first line start >= from  (from - 1 may be '\n' itself)
  while (p < e) {
    if (*p == '@') {
      const char *l2 = next_line(next_line(p, e), e);     // next-but-one line
      if (l2 < e && *l2 == '+') return p;
    }
    p = next_line(p, e);
  }
  return e;
}

void parse_fastq_range(const char *base, const char *b, const char *e_range, const char *e, Piece &out) {
  const char *p = fastq_record_start(base, b, e);
  while (p < e && p < e_range) {                          // records that START inside the range
    const char *seq = next_line(p, e);
    const char *seq_end = line_end(seq, e);
    size_t len = (size_t)(seq_end - seq);
    if (len && seq[len - 1] == '\r') --len;
    out.bases.append(seq, len);
    out.ends.push_back(out.bases.size());
    p = next_line(next_line(next_line(seq, e), e), e);    // behind sequence, '+' and quality lines
  }
}

// FASTA: [b, e_range) of the file; `k`: the last piece of a sequence that continues behind e_range gets k - 1 more bases
void parse_fasta_range(const char *base, const char *b, const char *e_range, const char *e, unsigned k, Piece &out) {
  const char *p = b;
  if (p != base && p[-1] != '\n') {                       // mid-line: is it a header line?  then skip it (it belongs to its record)
    const char *ls = p;
    while (ls > base && ls[-1] != '\n') --ls;
    if (*ls == '>') p = next_line(p, e);
  }
  bool open = false;                                      // a read is being filled
  while (p < e_range) {
    if (*p == '>' && (p == base || p[-1] == '\n')) {
      if (open) { out.ends.push_back(out.bases.size()); open = false; }
      p = next_line(p, e);
      continue;
    }
    const char *le = line_end(p, e);
    const char *stop = le < e_range ? le : e_range;       // never past the range: the next thread takes over there
    out.bases.append(p, (size_t)(stop - p));
    open = true;
    if (stop == le && le < e) { out.bases += '\n'; p = le + 1; } else p = stop;
  }
  if (open) {
    // the sequence may continue in the next range: repeat its next k - 1 bases (EOL characters do not count, a header ends it)
    unsigned need = k - 1;
    const char *q = p;
    while (need && q < e) {
      if (*q == '>' && (q == base || q[-1] == '\n')) break;
      if (*q != '\n' && *q != '\r') --need;
      out.bases += *q++;
    }
    out.ends.push_back(out.bases.size());
  }
}

void read_sequences(const std::string &path, unsigned k, std::string &bases, std::vector<uint64_t> &offsets) {
  if (offsets.empty()) offsets.push_back(0);
  const int fd = ::open(path.c_str(), O_RDONLY);
  if (fd < 0) { std::fprintf(stderr, "cannot open %s\n", path.c_str()); std::exit(1); }
  struct stat st;
  if (::fstat(fd, &st) != 0 || st.st_size == 0) { ::close(fd); return; }
  const size_t size = (size_t)st.st_size;
  void *map = ::mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
  ::close(fd);
  if (map == MAP_FAILED) { std::fprintf(stderr, "cannot map %s\n", path.c_str()); std::exit(1); }
  const char *base = (const char *)map, *e = base + size;
  const char *first = base;
  while (first < e && (*first == '\n' || *first == '\r')) ++first;
  const bool fastq = first < e && *first == '@';
  unsigned threads = std::max(1u, std::thread::hardware_concurrency());
  if (const char *ev = std::getenv("KH_READ_THREADS")) threads = std::max(1, std::atoi(ev));
  const size_t min_range = std::getenv("KH_READ_THREADS") ? 1 : (size_t)1 << 20;      // tests force tiny ranges
  threads = (unsigned)std::max<size_t>(1, std::min<size_t>(threads, size / min_range + 1));
  std::vector<Piece> pieces(threads);
  std::vector<std::thread> th;
  for (unsigned t = 0; t < threads; ++t) th.emplace_back([&, t]() {
    const char *b = base + size * t / threads, *er = base + size * (t + 1) / threads;
    if (fastq) parse_fastq_range(base, b, er, e, pieces[t]);
    else parse_fasta_range(base, b, er, e, k, pieces[t]);
  });
  for (auto &x : th) x.join();
  size_t total = bases.size();
  for (auto &pc : pieces) total += pc.bases.size();
  bases.reserve(total);
  for (auto &pc : pieces) {
    const uint64_t at = bases.size();
    bases += pc.bases;
    for (uint64_t end : pc.ends) offsets.push_back(at + end);
  }
  ::munmap(map, size);
}

// uniform random genome (mt19937_64 seed 1), reads from either strand (seed 2),
// 1 % substitutions (seed 3), no N
void synthetic_reads(const Options &o, std::string &bases, std::vector<uint64_t> &offsets) {
  static const char A[4] = {'A', 'C', 'G', 'T'};
  std::mt19937_64 g1(1), g2(2), g3(3);
  std::string genome(o.syn_genome, 'A');
  for (size_t i = 0; i < o.syn_genome; i += 32) {
    uint64_t r = g1();
    for (size_t j = i; j < std::min(i + 32, o.syn_genome); ++j, r >>= 2) genome[j] = A[r & 3];
  }
  bases.resize(o.syn_reads * o.syn_len);
  offsets.resize(o.syn_reads + 1);
  std::uniform_int_distribution<size_t> pos(0, o.syn_genome - o.syn_len);
  const uint64_t err_thresh = (uint64_t)(0.01 * 18446744073709551615.0);
  for (size_t r = 0; r < o.syn_reads; ++r) {
    const size_t s = pos(g2);
    const bool rc = g2() & 1;
    char *dst = &bases[r * o.syn_len];
    for (size_t j = 0; j < o.syn_len; ++j) {
      char c = rc ? genome[s + o.syn_len - 1 - j] : genome[s + j];
      int code = c == 'A' ? 0 : (c == 'C' ? 1 : (c == 'G' ? 2 : 3));
      if (rc) code = 3 - code;
      if (g3() < err_thresh) code = (code + 1 + (int)(g3() % 3)) & 3;
      dst[j] = A[code];
    }
    offsets[r] = r * o.syn_len;
  }
  offsets[o.syn_reads] = o.syn_reads * o.syn_len;
}

Options parse(int argc, char **argv) {
  Options o;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto next = [&]() -> const char * {
      if (i + 1 >= argc) { std::fprintf(stderr, "missing value for %s\n", a.c_str()); std::exit(2); }
      return argv[++i];
    };
    if (a == "-O" || a == "--out_filename") o.out = next();
    else if (a == "-G" || a == "--gpus") o.gpus = std::atoi(next());
    else if (a == "-A" || a == "--algo" || a == "-B" || a == "--output_algo") next();   // reader / writer ids: one path here
    else if (a == "--sample") o.sample = std::strtoull(next(), nullptr, 10);
    else if (a == "--reads-only") o.reads_only = next();
    else if (a == "--synthetic") {
      if (std::sscanf(next(), "%zu,%zu,%zu", &o.syn_reads, &o.syn_len, &o.syn_genome) < 1) { std::fprintf(stderr, "--synthetic reads[,len[,genome]]\n"); std::exit(2); }
    } else if (a == "-h" || a == "--help") {
      std::puts("BenchmarkKmerCounter [-G gpus] [-O out] [--sample n] [--synthetic reads[,len[,genome]]] [files...]");
      std::exit(0);
    } else if (a[0] == '-') { std::fprintf(stderr, "unknown option %s\n", a.c_str()); std::exit(2); }
    else o.files.push_back(a);
  }
  if (o.files.empty() && o.syn_reads == 0) o.syn_reads = 1000000;
  return o;
}

}  // namespace

int main(int argc, char **argv) {
  const Options o = parse(argc, argv);
  try {
    PhaseTimer tm;
    std::string bases;
    std::vector<uint64_t> offsets;

    tm.start();
    if (o.syn_reads) synthetic_reads(o, bases, offsets);
    for (auto &f : o.files) read_sequences(f, K, bases, offsets);
    if (offsets.empty()) offsets.push_back(0);
    tm.end("read", offsets.size() - 1);
    if (!o.reads_only.empty()) {
      std::FILE *f = std::fopen(o.reads_only.c_str(), "wb");
      if (!f) { std::fprintf(stderr, "cannot write %s\n", o.reads_only.c_str()); return 1; }
      const uint64_t nr = offsets.size() - 1;
      std::fwrite(&nr, 8, 1, f);
      std::fwrite(offsets.data(), 8, offsets.size(), f);
      std::fwrite(bases.data(), 1, bases.size(), f);
      std::fclose(f);
      return 0;
    }

    // KmerParser + lex_less (canonical)
    tm.start();
    std::vector<KmerT> kmers = khb::kmers_from_reads<K>(bases.data(), offsets, true);
    tm.end("parse", kmers.size());
    const uint64_t *kw = reinterpret_cast<const uint64_t *>(kmers.data());
    const size_t ns = std::min(o.sample, kmers.size());

    std::vector<uint64_t> out_keys;
    std::vector<uint32_t> out_vals;
    size_t n_found = 0, n_present = 0;
    if (o.gpus <= 1) {
      khb::hashmap_robinhood_offsets_count<KmerT, uint32_t, khb::hash::murmur3avx64> map;
      tm.start();
      map.insert(kmers.data(), kmers.data() + kmers.size(), 1u);
      tm.end("insert", map.size());

      std::vector<uint8_t> present(ns);
      tm.start();
      n_present = map.count(present.data(), kmers.data(), kmers.data() + ns);
      tm.end("count", n_present);

      std::vector<uint32_t> vals(ns);
      tm.start();
      n_found = map.find(vals.data(), kmers.data(), kmers.data() + ns, 0u);
      tm.end("find", n_found);

      tm.start();
      auto all = map.to_vector();
      out_keys.resize(all.size()); out_vals.resize(all.size());
      for (size_t i = 0; i < all.size(); ++i) { out_keys[i] = all[i].first.data[0]; out_vals[i] = all[i].second; }
      tm.end("export", all.size());
    } else {
      khd_map *m = khd_create(o.gpus, 1, KH_HASH_MURMUR64AVX, KH_HASH_MURMUR64AVX, KH_REDUCE_COUNT, 0);
      if (!m) { std::fprintf(stderr, "khd_create: %s\n", kh_last_error()); return 1; }
      tm.start();
      khb::detail::ok(khd_insert_host(m, kw, kmers.size(), 1u), "khd_insert");
      tm.end("insert", khd_size(m));

      std::vector<uint8_t> present(ns);
      tm.start();
      n_present = (size_t)khb::detail::ok(khd_count_host(m, kw, ns, present.data()), "khd_count");
      tm.end("count", n_present);

      std::vector<uint32_t> vals(ns);
      tm.start();
      n_found = (size_t)khb::detail::ok(khd_find_host(m, kw, ns, vals.data(), 0u), "khd_find");
      tm.end("find", n_found);

      tm.start();
      for (int r = 0; r < o.gpus; ++r) {
        const size_t ls = khd_local_size(m, r), at = out_keys.size();
        out_keys.resize(at + ls); out_vals.resize(at + ls);
        khb::detail::ok(khd_export_rank_host(m, r, out_keys.data() + at, out_vals.data() + at, ls), "khd_export");
        std::printf("rank %d local size %zu\n", r, ls);
      }
      tm.end("export", out_keys.size());
      khd_destroy(m);
    }

    // [Kmer][count] records, CountType = uint16_t (benchmark/BenchmarkKmerCounter.cpp:184,1802-2050)
    if (!o.out.empty()) {
      tm.start();
      std::FILE *f = std::fopen(o.out.c_str(), "wb");
      if (!f) { std::fprintf(stderr, "cannot write %s\n", o.out.c_str()); return 1; }
      std::vector<unsigned char> rec(out_keys.size() * 10);
      for (size_t i = 0; i < out_keys.size(); ++i) {
        const CountType c = (CountType)out_vals[i];
        std::memcpy(&rec[i * 10], &out_keys[i], 8);
        std::memcpy(&rec[i * 10 + 8], &c, 2);
      }
      std::fwrite(rec.data(), 1, rec.size(), f);
      std::fclose(f);
      tm.end("write", out_keys.size());
    }

    uint64_t total = 0;
    for (uint32_t v : out_vals) total += v;
    std::printf("reads %zu kmers %zu distinct %zu total_count %llu sample_present %zu sample_found %zu\n", offsets.size() - 1,
                kmers.size(), out_keys.size(), (unsigned long long)total, n_present, n_found);
    tm.report("kmercounter");
    return total == kmers.size() ? 0 : 3;
  } catch (const std::exception &e) {
    std::fprintf(stderr, "error: %s\n", e.what());
    return 1;
  }
}
