// kmer_api.cu -- k-mer generation and canonicalisation kernels + their C ABI.
//
// Replaces:
//   KmerParser::operator()            KI/io/kmer_parser.hpp:232-293     all windows of a sequence
//   NSplitSequencesIterator           KI/io/filtered_sequence_iterator.hpp:225-247,429-440  cut at N/n
//   EOL filter                        KI/utils/file_utils.hpp:43-47     '\n' '\r' dropped
//   Kmer::nextFromChar                KI/common/kmer.hpp:716-726        kmer = (kmer << 2 | code) & mask
//   reverse_complement + lex_less     KI/common/kmer.hpp:1078-1087, KI/common/kmer_transform.hpp:109-125
//
// The reference shifts one base at a time through a serial dependency.  Here one warp owns one
// read and eats 32 bytes per step: the lanes classify their byte, the clean (non-EOL) 2-bit codes
// of the step are packed into one little-endian 64-bit word with two warp REDUX-OR operations, and
// every lane cuts its own window out of {previous 32*W bases, this step's bases} with funnel
// shifts -- no lane waits for another lane's k-mer.  In that little-endian window X the reverse
// complement is simply ~X (masked), and the forward k-mer is the 2-bit-group reversal of X.
// Two passes (count per read -> exclusive scan -> emit) keep the reference's output order.
#include <algorithm>
#include <cub/device/device_scan.cuh>
#include "kh_common.h"
#include "kh_kmer.cuh"
#include "kh_table.cuh"
#include "kh_kmergen.cuh"

namespace kh {

// shift the little-endian multiword value v[0..NV) right by `bits` (0 <= bits < 64*NV) and keep W words
template <int W>
__device__ __forceinline__ void extract_window(uint64_t v0, uint64_t v1, uint64_t v2, int bits, uint64_t (&x)[W]) {
  if (bits >= 64) { v0 = v1; v1 = v2; v2 = 0; bits -= 64; }
  if (bits >= 64) { v0 = v1; v1 = v2; v2 = 0; bits -= 64; }
  if (W == 1) {
    x[0] = bits ? ((v0 >> bits) | (v1 << (64 - bits))) : v0;
  } else {
    x[0] = bits ? ((v0 >> bits) | (v1 << (64 - bits))) : v0;
    x[W - 1] = bits ? ((v1 >> bits) | (v2 << (64 - bits))) : v1;
  }
}

// One warp per read.  EMIT = false: count the k-mers of every read (counts[r]).
// EMIT = true: write them at out_keys[out_off[r] ...].
template <int W, bool EMIT>
__global__ void __launch_bounds__(256) kmers_from_reads_kernel(const unsigned char *bases, const uint64_t *offsets,
                                                               uint64_t n_reads, int k, int canonical,
                                                               uint64_t *counts, const uint64_t *out_off,
                                                               uint64_t *out_keys, uint64_t out_cap) {
  const int lane = threadIdx.x & 31;
  const uint32_t lt = (1u << lane) - 1u;
  const uint64_t warps_total = (uint64_t)gridDim.x * (blockDim.x >> 5);
  const uint64_t warp_id = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  constexpr int PREV = 32 * W;                 // bases of history kept (>= k - 1)
  // mask of the 2k significant bits in the top word
  const int topbits = 2 * k - 64 * (W - 1);
  const uint64_t topmask = topbits >= 64 ? ~0ULL : ((1ULL << topbits) - 1ULL);

  for (uint64_t r = warp_id; r < n_reads; r += warps_total) {
    const uint64_t beg = offsets[r], end = offsets[r + 1];
    uint64_t pv0 = 0, pv1 = 0;                 // previous PREV bases, little-endian, newest at the top
    uint64_t pn = ~0ULL;                       // invalid-flags of the previous 64 positions (1 = N or "before the read")
    uint64_t emitted = 0;
    const uint64_t obase = EMIT ? out_off[r] : 0;
    for (uint64_t p0 = beg; p0 < end; p0 += 32) {
      const uint64_t pos = p0 + lane;
      const unsigned char c = pos < end ? bases[pos] : (unsigned char)'\n';   // padding behaves like EOL
      const bool skip = (c == '\n') || (c == '\r');
      const bool isn = (c == 'N') || (c == 'n');
      const uint32_t clean = __ballot_sync(FULL_MASK, !skip);
      const int ci = __popc(clean & lt);       // index of this lane's base among the step's clean bases
      const int cnt = __popc(clean);
      const uint32_t code = skip ? 0u : dna_code(c);
      // pack: clean base ci -> bits [2ci, 2ci+2)
      const uint32_t lo = __reduce_or_sync(FULL_MASK, (!skip && ci < 16) ? (code << (2 * ci)) : 0u);
      const uint32_t hi = __reduce_or_sync(FULL_MASK, (!skip && ci >= 16) ? (code << (2 * (ci - 16))) : 0u);
      const uint32_t nm = __reduce_or_sync(FULL_MASK, (!skip && isn) ? (1u << ci) : 0u);
      const uint64_t cur = ((uint64_t)hi << 32) | lo;

      // window of this lane: clean positions [PREV + ci - k + 1, PREV + ci] of {prev, cur}
      const int start = PREV + ci - k + 1;     // >= 1 because k <= 32 W
      uint64_t x[W];
      if (W == 1) extract_window<W>(pv0, cur, 0, 2 * start, x);
      else extract_window<W>(pv0, pv1, cur, 2 * start, x);
      x[W - 1] &= topmask;
      // validity: no invalid flag among the k positions.  flags: {pn (64 older positions), nm (this step)}
      // position PREV + j of the value stream corresponds to flag bit 64 + j.
      const int fstart = 64 + ci - k + 1;      // in [1, 95]
      uint64_t f0 = pn, f1 = (uint64_t)nm;
      int fb = fstart;
      if (fb >= 64) { f0 = f1; f1 = 0; fb -= 64; }
      uint64_t fw = fb ? ((f0 >> fb) | (f1 << (64 - fb))) : f0;
      const uint64_t kmask = k >= 64 ? ~0ULL : ((1ULL << k) - 1ULL);
      const bool valid = !skip && ((fw & kmask) == 0);
      const uint32_t vm = __ballot_sync(FULL_MASK, valid);
      if (EMIT && valid) {
        const uint64_t o = obase + emitted + __popc(vm & lt);
        if (o < out_cap) {
          // x is the window little-endian (oldest base lowest): revcomp = ~x masked; forward = reversed groups
          uint64_t rc[W], fwd[W];
#pragma unroll
          for (int q = 0; q < W; ++q) rc[q] = ~x[q];
          rc[W - 1] &= topmask;
          // forward = rev2bit over the whole W-word value, shifted down by 64W - 2k
          const int sh = 64 * W - 2 * k;
          if (W == 1) fwd[0] = rev2bit64(x[0]) >> sh;
          else {
            const uint64_t t0 = rev2bit64(x[W - 1]), t1 = rev2bit64(x[0]);
            fwd[0] = sh ? ((t0 >> sh) | (t1 << (64 - sh))) : t0;
            fwd[W - 1] = t1 >> sh;
          }
          const bool use_rc = canonical && !kmer_less<W>(fwd, rc);
#pragma unroll
          for (int q = 0; q < W; ++q) out_keys[o * W + q] = use_rc ? rc[q] : fwd[q];
        }
      }
      emitted += __popc(vm);
      // slide the history by cnt clean bases (0..32)
      if (cnt) {
        const int b2 = 2 * cnt;                // 2..64
        if (W == 1) {
          pv0 = (b2 == 64) ? cur : ((pv0 >> b2) | (cur << (64 - b2)));
        } else {
          if (b2 == 64) { pv0 = pv1; pv1 = cur; }
          else { pv0 = (pv0 >> b2) | (pv1 << (64 - b2)); pv1 = (pv1 >> b2) | (cur << (64 - b2)); }
        }
        pn = (cnt == 32) ? ((pn >> 32) | ((uint64_t)nm << 32)) : ((pn >> cnt) | ((uint64_t)nm << (64 - cnt)));
      }
    }
    if (!EMIT && lane == 0) counts[r] = emitted;
  }
}

template <int W>
__global__ void canonicalize_kernel(const uint64_t *in, uint64_t n, int alpha, int k, int mode, uint64_t *out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t key[W], res[W];
    load_key<W>(in, i, key);
    if (mode == 0) kmer_canonical_alpha<W>(key, alpha, k, res); else kmer_revcomp_alpha<W>(key, alpha, k, res);
#pragma unroll
    for (int q = 0; q < W; ++q) out[i * W + q] = res[q];
  }
}

// one THREAD per read, serial: the fallback of the tile kernel for the other alphabets (EOL-dense input only; the restatement of
// KmerParser over the N-splitting, EOL-skipping iterator with Kmer::nextFromChar, nothing more)
template <int W, int B, bool EMIT>
__global__ void kmers_serial_kernel(const unsigned char *bases, const uint64_t *offsets, uint64_t n_reads, int k, int canonical,
                                    uint64_t *counts, const uint64_t *out_off, uint64_t *out_keys, uint64_t out_cap) {
  const int alpha = B == 2 ? KH_ALPHA_DNA : B == 3 ? KH_ALPHA_DNA5 : KH_ALPHA_DNA16;
  const int topbits = B * k - 64 * (W - 1);
  const uint64_t topmask = topbits >= 64 ? ~0ULL : ((1ULL << topbits) - 1ULL);
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; r < n_reads; r += stride) {
    uint64_t w[W];
#pragma unroll
    for (int q = 0; q < W; ++q) w[q] = 0;
    uint64_t run = 0, emitted = 0;
    const uint64_t obase = EMIT ? out_off[r] : 0;
    for (uint64_t p = offsets[r]; p < offsets[r + 1]; ++p) {
      const unsigned char c = bases[p];
      if (c == 'N' || c == 'n') { run = 0; continue; }
      if (c == '\n' || c == '\r') continue;
      if (W == 2) w[W - 1] = ((w[W - 1] << B) | (w[0] >> (64 - B)));
      w[0] = (w[0] << B) | alpha_code<B>(c);
      w[W - 1] &= topmask;
      if (++run >= (uint64_t)k) {
        if (EMIT && obase + emitted < out_cap) {
          uint64_t res[W];
          if (canonical) kmer_canonical_alpha<W>(w, alpha, k, res);
          else {
#pragma unroll
            for (int q = 0; q < W; ++q) res[q] = w[q];
          }
#pragma unroll
          for (int q = 0; q < W; ++q) out_keys[(obase + emitted) * W + q] = res[q];
        }
        ++emitted;
      }
    }
    if (!EMIT) counts[r] = emitted;
  }
}

// the warp-per-read kernel: count per read -> exclusive scan -> emit.  No bound on EOL characters inside a window; the
// fallback of the tile kernel and the reference point of its tests (KH_KMERGEN=0).
int64_t kmers_from_reads_warp(int device, const char *d_bases, const uint64_t *d_offsets, uint64_t n_reads, int k,
                              int canonical, uint64_t *d_out_keys, uint64_t out_cap) {
  const int W = kmer_words(k);
  const int sms = sm_count(device);
  const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n_reads + 7) / 8, (uint64_t)sms * 8));
  DevBuf<uint64_t> counts(n_reads + 1), offs(n_reads + 1);
  const unsigned char *b = reinterpret_cast<const unsigned char *>(d_bases);
  KH_CUDA(cudaMemsetAsync(counts.p + n_reads, 0, sizeof(uint64_t), 0));
  if (W == 1) KH_LAUNCH((kmers_from_reads_kernel<1, false>), grid, 256, 0, 0, b, d_offsets, n_reads, k, canonical, counts.p, nullptr, nullptr, 0);
  else KH_LAUNCH((kmers_from_reads_kernel<2, false>), grid, 256, 0, 0, b, d_offsets, n_reads, k, canonical, counts.p, nullptr, nullptr, 0);
  size_t tmp_bytes = 0;
  KH_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts.p, offs.p, n_reads + 1, 0));
  DevBuf<unsigned char> tmp(tmp_bytes);
  KH_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, counts.p, offs.p, n_reads + 1, 0));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  uint64_t total = 0;
  KH_CUDA(cudaMemcpy(&total, offs.p + n_reads, sizeof(uint64_t), cudaMemcpyDeviceToHost));
  if (d_out_keys && total) {
    if (W == 1) KH_LAUNCH((kmers_from_reads_kernel<1, true>), grid, 256, 0, 0, b, d_offsets, n_reads, k, canonical, nullptr, offs.p, d_out_keys, out_cap);
    else KH_LAUNCH((kmers_from_reads_kernel<2, true>), grid, 256, 0, 0, b, d_offsets, n_reads, k, canonical, nullptr, offs.p, d_out_keys, out_cap);
    KH_CUDA(cudaStreamSynchronize(0));
  }
  return (int64_t)total;
}

template <int W, int B>
int64_t kmers_from_reads_serial(int device, const unsigned char *b, const uint64_t *d_offsets, uint64_t n_reads, int k,
                                int canonical, uint64_t *d_out_keys, uint64_t out_cap) {
  const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n_reads + 127) / 128, (uint64_t)sm_count(device) * 16));
  DevBuf<uint64_t> counts(n_reads + 1), offs(n_reads + 1);
  KH_CUDA(cudaMemsetAsync(counts.p + n_reads, 0, sizeof(uint64_t), 0));
  KH_LAUNCH((kmers_serial_kernel<W, B, false>), grid, 128, 0, 0, b, d_offsets, n_reads, k, canonical, counts.p, nullptr, nullptr, 0);
  size_t tmp_bytes = 0;
  KH_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts.p, offs.p, n_reads + 1, 0));
  DevBuf<unsigned char> tmp(tmp_bytes);
  KH_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, counts.p, offs.p, n_reads + 1, 0));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  uint64_t total = 0;
  KH_CUDA(cudaMemcpy(&total, offs.p + n_reads, sizeof(uint64_t), cudaMemcpyDeviceToHost));
  if (d_out_keys && total) {
    KH_LAUNCH((kmers_serial_kernel<W, B, true>), grid, 128, 0, 0, b, d_offsets, n_reads, k, canonical, nullptr, offs.p, d_out_keys, out_cap);
    KH_CUDA(cudaStreamSynchronize(0));
  }
  return (int64_t)total;
}

// the streaming tile kernel (kh_kmergen.cuh): one pass, output order kept by a decoupled look-back
int64_t kmers_from_reads_dev(int device, int alpha, const char *d_bases, const uint64_t *d_offsets, uint64_t n_reads, int k,
                             int canonical, uint64_t *d_out_keys, uint64_t out_cap) {
  const int B = alpha_bits(alpha);
  const int W = kmer_words_alpha(alpha, k);
  if (!B) { set_error("alphabet must be 4 (DNA), 5 (DNA5) or 16 (DNA16)"); return KH_ERR_INVALID; }
  if (k < 1 || W < 1 || W > 2) { set_error("k must be in [1,%d] for this alphabet", 128 / B); return KH_ERR_INVALID; }
  if (n_reads == 0) return 0;
  static const bool tiles_on = !(getenv("KH_KMERGEN") && atoi(getenv("KH_KMERGEN")) == 0);
  const unsigned char *ub = reinterpret_cast<const unsigned char *>(d_bases);
  auto slow = [&]() -> int64_t {
    if (B == 2) return kmers_from_reads_warp(device, d_bases, d_offsets, n_reads, k, canonical, d_out_keys, out_cap);
    if (B == 3) return W == 1 ? kmers_from_reads_serial<1, 3>(device, ub, d_offsets, n_reads, k, canonical, d_out_keys, out_cap)
                              : kmers_from_reads_serial<2, 3>(device, ub, d_offsets, n_reads, k, canonical, d_out_keys, out_cap);
    return W == 1 ? kmers_from_reads_serial<1, 4>(device, ub, d_offsets, n_reads, k, canonical, d_out_keys, out_cap)
                  : kmers_from_reads_serial<2, 4>(device, ub, d_offsets, n_reads, k, canonical, d_out_keys, out_cap);
  };
  if (!tiles_on) return slow();
  uint64_t h_off[2] = {0, 0};
  KH_CUDA(cudaMemcpy(&h_off[0], d_offsets, sizeof(uint64_t), cudaMemcpyDeviceToHost));
  KH_CUDA(cudaMemcpy(&h_off[1], d_offsets + n_reads, sizeof(uint64_t), cudaMemcpyDeviceToHost));
  if (h_off[1] <= h_off[0]) return 0;
  GenArgs a;
  a.delta = (uint64_t)(reinterpret_cast<uintptr_t>(d_bases) & 15u);
  a.bases = reinterpret_cast<const unsigned char *>(d_bases) - a.delta;
  a.offsets = d_offsets; a.n_reads = n_reads;
  a.first = h_off[0] + a.delta; a.end = h_off[1] + a.delta;
  a.k = k; a.canonical = canonical;
  a.n_tiles = (a.end + GEN_SPAN - 1) / GEN_SPAN;
  a.out_keys = (d_out_keys && out_cap) ? d_out_keys : nullptr; a.out_cap = out_cap;
  DevBuf<unsigned long long> state(a.n_tiles);
  DevBuf<uint64_t> tile_read(a.n_tiles);
  DevBuf<int> flag(1);
  KH_CUDA(cudaMemsetAsync(state.p, 0, a.n_tiles * sizeof(unsigned long long), 0));
  KH_CUDA(cudaMemsetAsync(flag.p, 0, sizeof(int), 0));
  KH_LAUNCH(gen_tile_reads_kernel, (int)((a.n_tiles + 255) / 256), 256, 0, 0, d_offsets, n_reads, a.delta, a.n_tiles, tile_read.p);
  a.state = state.p; a.tile_read = tile_read.p; a.fallback = flag.p;
  const int sms = sm_count(device);
  const size_t smem = W == 1 ? gen_smem_bytes<1>() : gen_smem_bytes<2>();
  const int per_sm = std::max<int>(1, std::min<int>(2, (int)((227 * 1024) / (smem + 1024))));
  const int grid = (int)std::min<uint64_t>(a.n_tiles, (uint64_t)sms * per_sm);
#define KH_GEN_LAUNCH(Wv, Bv)                                                                                                  \
  do {                                                                                                                         \
    KH_CUDA(cudaFuncSetAttribute((kmer_tiles_kernel<Wv, false, Bv>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    KH_LAUNCH((kmer_tiles_kernel<Wv, false, Bv>), grid, GEN_THREADS, smem, 0, a);                                               \
  } while (0)
  if (B == 2) { if (W == 1) KH_GEN_LAUNCH(1, 2); else KH_GEN_LAUNCH(2, 2); }
  else if (B == 3) { if (W == 1) KH_GEN_LAUNCH(1, 3); else KH_GEN_LAUNCH(2, 3); }
  else { if (W == 1) KH_GEN_LAUNCH(1, 4); else KH_GEN_LAUNCH(2, 4); }
#undef KH_GEN_LAUNCH
  unsigned long long last = 0;
  int h_flag = 0;
  KH_CUDA(cudaMemcpyAsync(&last, state.p + a.n_tiles - 1, sizeof(last), cudaMemcpyDeviceToHost, 0));
  KH_CUDA(cudaMemcpyAsync(&h_flag, flag.p, sizeof(int), cudaMemcpyDeviceToHost, 0));
  KH_CUDA(cudaStreamSynchronize(0));
  if (h_flag) return slow();                                         // EOL-dense input
  return (int64_t)(last & ~GEN_ST_MASK);
}

}  // namespace kh

using namespace kh;

extern "C" {

int kh_kmer_words_alpha(int alphabet, int k) { return (k >= 1 && alpha_bits(alphabet)) ? kmer_words_alpha(alphabet, k) : 0; }

int64_t kh_kmers_from_reads_alpha_dev(int device, int alphabet, const char *d_bases, const uint64_t *d_offsets, uint64_t n_reads,
                                      int k, int canonical, uint64_t *d_out_keys, uint64_t out_cap) {
  return guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    DeviceGuard g(device);
    return kmers_from_reads_dev(device, alphabet, d_bases, d_offsets, n_reads, k, canonical, d_out_keys, out_cap);
  });
}
int64_t kh_kmers_from_reads_dev(int device, const char *d_bases, const uint64_t *d_offsets, uint64_t n_reads,
                                int k, int canonical, uint64_t *d_out_keys, uint64_t out_cap) {
  return kh_kmers_from_reads_alpha_dev(device, KH_ALPHA_DNA, d_bases, d_offsets, n_reads, k, canonical, d_out_keys, out_cap);
}

int64_t kh_kmers_from_reads_alpha_host(int device, int alphabet, const char *bases, const uint64_t *offsets, uint64_t n_reads,
                                       int k, int canonical, uint64_t *out_keys, uint64_t out_cap) {
  return guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    const int W = kh_kmer_words_alpha(alphabet, k);
    if (W < 1 || W > 2) { set_error("alphabet 4 / 5 / 16 with k up to 64 / 42 / 32"); return KH_ERR_INVALID; }
    if (n_reads == 0) return 0;
    DeviceGuard g(device);
    const uint64_t nbytes = offsets[n_reads];
    DevBuf<char> db(std::max<uint64_t>(nbytes, 1));
    DevBuf<uint64_t> doff(n_reads + 1);
    KH_CUDA(cudaMemcpy(db.p, bases, nbytes, cudaMemcpyHostToDevice));
    KH_CUDA(cudaMemcpy(doff.p, offsets, (n_reads + 1) * 8, cudaMemcpyHostToDevice));
    DevBuf<uint64_t> dout;
    if (out_keys && out_cap) dout.alloc(out_cap * W);
    const int64_t total = kmers_from_reads_dev(device, alphabet, db.p, doff.p, n_reads, k, canonical, dout.p, out_keys ? out_cap : 0);
    if (total > 0 && out_keys) {
      const uint64_t c = std::min<uint64_t>((uint64_t)total, out_cap);
      KH_CUDA(cudaMemcpy(out_keys, dout.p, c * W * 8, cudaMemcpyDeviceToHost));
    }
    return total;
  });
}
int64_t kh_kmers_from_reads_host(int device, const char *bases, const uint64_t *offsets, uint64_t n_reads,
                                 int k, int canonical, uint64_t *out_keys, uint64_t out_cap) {
  return kh_kmers_from_reads_alpha_host(device, KH_ALPHA_DNA, bases, offsets, n_reads, k, canonical, out_keys, out_cap);
}

static int canon_impl(int device, int alpha, int k, const uint64_t *d_in, uint64_t n, uint64_t *d_out, int mode) {
  const int W = kh_kmer_words_alpha(alpha, k);
  if (W < 1 || W > 2) { set_error("alphabet 4 / 5 / 16 with k up to 64 / 42 / 32"); return KH_ERR_INVALID; }
  if (n == 0) return KH_OK;
  const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count(device) * 8));
  if (W == 1) KH_LAUNCH((canonicalize_kernel<1>), grid, 256, 0, 0, d_in, n, alpha, k, mode, d_out);
  else KH_LAUNCH((canonicalize_kernel<2>), grid, 256, 0, 0, d_in, n, alpha, k, mode, d_out);
  KH_CUDA(cudaStreamSynchronize(0));
  return KH_OK;
}

int kh_canonicalize_dev(int device, int k, const uint64_t *d_in, uint64_t n, uint64_t *d_out) {
  return (int)guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    DeviceGuard g(device);
    return canon_impl(device, KH_ALPHA_DNA, k, d_in, n, d_out, 0);
  });
}

static int canon_host(int device, int alpha, int k, const uint64_t *in, uint64_t n, uint64_t *out, int mode) {
  return (int)guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    const int W = kh_kmer_words_alpha(alpha, k);
    if (W < 1 || W > 2) { set_error("alphabet 4 / 5 / 16 with k up to 64 / 42 / 32"); return KH_ERR_INVALID; }
    if (n == 0) return KH_OK;
    DeviceGuard g(device);
    DevBuf<uint64_t> d(n * W);
    KH_CUDA(cudaMemcpy(d.p, in, n * W * 8, cudaMemcpyHostToDevice));
    const int rc = canon_impl(device, alpha, k, d.p, n, d.p, mode);
    if (rc != KH_OK) return rc;
    KH_CUDA(cudaMemcpy(out, d.p, n * W * 8, cudaMemcpyDeviceToHost));
    return KH_OK;
  });
}
int kh_canonicalize_host(int device, int k, const uint64_t *in, uint64_t n, uint64_t *out) { return canon_host(device, KH_ALPHA_DNA, k, in, n, out, 0); }
int kh_revcomp_host(int device, int k, const uint64_t *in, uint64_t n, uint64_t *out) { return canon_host(device, KH_ALPHA_DNA, k, in, n, out, 1); }
int kh_canonicalize_alpha_host(int device, int alphabet, int k, const uint64_t *in, uint64_t n, uint64_t *out) { return canon_host(device, alphabet, k, in, n, out, 0); }
int kh_revcomp_alpha_host(int device, int alphabet, int k, const uint64_t *in, uint64_t n, uint64_t *out) { return canon_host(device, alphabet, k, in, n, out, 1); }

}  // extern "C"
