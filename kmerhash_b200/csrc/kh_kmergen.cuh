// kh_kmergen.cuh -- block-level streaming k-mer generation (2-bit DNA, k <= 64), one pass over the base buffer.
//
// Replaces the hot loop of KmerParser::operator() (KI/io/kmer_parser.hpp:232-293) over the N-splitting, EOL-skipping
// sequence iterator (KI/io/filtered_sequence_iterator.hpp:225-247,429-440; KI/utils/file_utils.hpp:43-47) with
// Kmer::nextFromChar (KI/common/kmer.hpp:716-726) and, fused, reverse_complement + lex_less (KI/common/kmer.hpp:1078-1087,
// KI/common/kmer_transform.hpp:109-125).  The reference pushes one base at a time through a serial shift per read; the
// first GPU version gave one warp per read.  Here the unit of work is a TILE of the byte buffer, independent of the reads:
//
//   1. a CTA loads GEN_TILE_BYTES consecutive bytes (16 per thread, one coalesced 16-byte load): GEN_LOOKBACK bytes of
//      context + GEN_SPAN bytes whose k-mers (by END position) this tile emits;
//   2. every thread classifies its bytes (EOL / N / 2-bit code) and packs its clean bases; a block scan of the clean counts
//      gives every thread its place in the tile's COMPACT base sequence (EOL characters squeezed out), where it ORs its
//      packed codes and N flags into shared-memory bit arrays; the read starts that fall into the tile (binary search in
//      `offsets`, then one thread per start) become bits of a third array;
//   3. every thread takes 16 consecutive compact positions: validity of the 16 windows from the bit arrays, a block scan of
//      the valid counts, then a rolling update  fwd = (fwd << 2 | c) & mask,  rc = rc >> 2 | (3 - c) << (2k - 2)  from a
//      start window cut out of the packed codes with funnel shifts -- no per-k-mer gather, no warp reduction per step;
//   4. the tile's k-mers go to a shared-memory tile in output order and leave with coalesced 16-byte stores.  Where the tile's
//      output begins is settled without a second pass by a decoupled look-back over per-tile aggregates (the grid is
//      persistent and takes the tiles round-robin, so a tile only ever waits for tiles of CTAs that are resident and run).
// The first read of every tile comes from a tiny pre-kernel (one binary search per tile, all in parallel); the next tile's
// bytes are fetched into registers while the current tile is processed.
//
// Output order = the reference's: reads in order, windows in order.  A window that would reach behind the tile's context
// (more than GEN_LOOKBACK - k EOL characters inside k bases) cannot be decided here: the kernel raises `fallback` and the
// host repeats the call with the warp-per-read kernel (kmer_api.cu), which has no such bound.
#pragma once
#include <cstdint>
#include <cub/block/block_scan.cuh>
#include "kh_kmer.cuh"
#include "kh_table.cuh"
#include "kh_window.cuh"

namespace kh {

constexpr int GEN_THREADS = 512;
constexpr int GEN_TILE_BYTES = GEN_THREADS * 16;          // 8192
constexpr int GEN_LOOKBACK = 512;
constexpr int GEN_SPAN = GEN_TILE_BYTES - GEN_LOOKBACK;   // 7680 output positions per tile = 15 warps x 32 lanes x 16: the last warp
                                                          // is free for the look-back while the others generate
constexpr int GEN_PAD = 64;                               // compact positions of zero history in front of the tile
constexpr int GEN_RUN = 16;                               // compact positions per thread in the rolling stage
constexpr int GEN_BITS = GEN_TILE_BYTES + GEN_PAD + 64;   // compact positions the bit arrays hold (incl. slack for 64-bit reads)
constexpr int GEN_PC_WORDS = GEN_BITS / 8;                // packed codes: up to 4 bits per position (DNA 2, DNA5 3, DNA16 4)

constexpr unsigned long long GEN_ST_AGG = 1ULL << 62, GEN_ST_INC = 2ULL << 62, GEN_ST_MASK = 3ULL << 62;

struct GenArgs {
  const unsigned char *bases;      // 16-byte aligned origin (the caller's pointer rounded down; `delta` = the difference)
  const uint64_t *offsets;         // [n_reads + 1], relative to the caller's pointer
  uint64_t n_reads, delta;
  uint64_t first, end;             // offsets[0] + delta, offsets[n_reads] + delta
  int k, canonical;
  uint64_t n_tiles;
  const uint64_t *tile_read;       // [n_tiles] first read whose start lies at or behind the tile's first byte (gen_tile_reads_kernel)
  unsigned long long *state;       // [n_tiles] look-back words
  uint64_t *out_keys;              // may be null: count only
  uint64_t out_cap;
  int *fallback;
  // SCATTER mode (fused with the multi-GPU sender, exchange.cu): the tile's k-mers do not go to out_keys but straight to their
  // OWNER's slab: dist hash -> owner, tile sorted by owner in shared memory, runs stored through peer pointers (NVLink).
  // Output order is free there (COUNT commutes), so no look-back: one global atomicAdd per (tile, owner) on `cursors`.
  BinSpec dist;
  BinDst dst;
  unsigned long long slab_cap;
  unsigned long long *cursors;     // [p]: ends as the number of k-mers this rank has for every owner
};

template <int W> constexpr size_t gen_smem_bytes(bool scatter = false) {
  // out tile | packed codes (u32) | N bits | start bits | per-thread clean mask + compact base | SCATTER: sorted index of the tile
  return (size_t)GEN_SPAN * 8 * W + (size_t)GEN_PC_WORDS * 4 + 2 * (size_t)(GEN_BITS / 32) * 4 + (size_t)GEN_THREADS * 8 + 64 +
         (scatter ? (size_t)GEN_SPAN * 2 : 0);
}

// 64 bits starting at bit position `bit` of a little-endian u32 array
__device__ __forceinline__ uint64_t bits64(const uint32_t *a, uint32_t bit) {
  const uint32_t w = bit >> 5, s = bit & 31u;
  const uint32_t x0 = a[w], x1 = a[w + 1], x2 = a[w + 2];
  const uint32_t lo = __funnelshift_r(x0, x1, s), hi = __funnelshift_r(x1, x2, s);
  return ((uint64_t)hi << 32) | lo;
}

// first read whose start lies at or behind byte max(lo, 0) of every tile: lower bound over offsets + delta
static __global__ void gen_tile_reads_kernel(const uint64_t *offsets, uint64_t n_reads, uint64_t delta, uint64_t n_tiles, uint64_t *tile_read) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  const int64_t lo = (int64_t)(t * GEN_SPAN) - GEN_LOOKBACK;
  const uint64_t want = lo > 0 ? (uint64_t)lo : 0ULL;
  uint64_t l = 0, h = n_reads + 1;
  while (l < h) {
    const uint64_t mid = (l + h) >> 1;
    if (offsets[mid] + delta < want) l = mid + 1; else h = mid;
  }
  tile_read[t] = l;
}

// Where k-mer number i of a tile lies in the shared-memory out tile.  Every thread emits a run of up to 16 consecutive k-mers, so
// in a plain layout the 32 lanes of a store instruction would hit addresses 128 bytes (W = 2: 256) apart: one bank, 32-way
// conflict.  XOR of the low four index bits with the next four spreads them; the layout is a bijection inside every aligned group
// of 16, so the copy-out can still walk the tile in order.
__device__ __forceinline__ uint32_t gen_slot(uint32_t i) { return i ^ ((i >> 4) & 15u); }

// four bytes at a time: bit 7 of every byte that equals the constant (exact: the add cannot carry across bytes)
__device__ __forceinline__ uint32_t eq4(uint32_t w, uint32_t c4) {
  const uint32_t z = w ^ c4;
  return ~(((z & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | z | 0x7F7F7F7Fu);
}
// one flag bit per byte (bit 7 of every byte of m) -> bits 0,1,2,3 / bits 0,2,4,6 of the result
__device__ __forceinline__ uint32_t mask4(uint32_t m) { return ((m >> 7) * 0x01020408u) >> 24; }
__device__ __forceinline__ uint32_t spread4(uint32_t m) { return ((m >> 7) * 0x01041040u) >> 24; }

// Consumer: what happens to a tile's k-mers.  store(): coalesced copy of the shared-memory tile to out_keys (the stand-alone API).
// B = bits per base: 2 (DNA), 3 (DNA5), 4 (DNA16).  The DNA path classifies four bytes per step and cuts its start window out of
// the packed codes; the other alphabets go through their code function byte by byte and roll the start window up from its
// k - 1 bases (they are the reference's side configurations, benchmark/CMakeLists.txt:175-196, not the measured path).
template <int W, bool SCATTER, int B = 2>
__global__ void __launch_bounds__(GEN_THREADS, 2) kmer_tiles_kernel(const __grid_constant__ GenArgs a) {
  using Scan = cub::BlockScan<uint32_t, GEN_THREADS>;
  __shared__ typename Scan::TempStorage scan_tmp;
  __shared__ unsigned long long s_base;
  __shared__ uint32_t s_firstbreak;
  __shared__ uint32_t s_cnt[WIN_MAX_DST], s_lo[WIN_MAX_DST + 1], s_vlo[WIN_MAX_DST + 1];
  __shared__ unsigned long long s_gbase[WIN_MAX_DST];
  extern __shared__ __align__(16) unsigned char gen_smem[];
  uint64_t *otile = reinterpret_cast<uint64_t *>(gen_smem);
  uint32_t *pc = reinterpret_cast<uint32_t *>(otile + (size_t)GEN_SPAN * W);   // 2 bits per compact position
  uint32_t *nb = pc + GEN_PC_WORDS;                                             // 1 bit: N
  uint32_t *sb = nb + GEN_BITS / 32;                                            // 1 bit: a read starts here
  uint32_t *cbase = sb + GEN_BITS / 32;                                         // [GEN_THREADS] compact index of the thread's first byte
  uint32_t *cmask = cbase + GEN_THREADS;                                        // [GEN_THREADS] clean mask of the thread's 16 bytes
  uint16_t *sidx = reinterpret_cast<uint16_t *>(cmask + GEN_THREADS + 16);      // SCATTER: [GEN_SPAN] tile position of every sorted k-mer
  const int tid = threadIdx.x, lane = tid & 31;
  const int k = a.k;
  const int topbits = B * k - 64 * (W - 1);
  const uint64_t topmask = topbits >= 64 ? ~0ULL : ((1ULL << topbits) - 1ULL);

  // the 16 bytes of this thread in tile t; bytes outside [first, end) read as EOL
  auto fetch = [&](uint64_t t) -> uint4 {
    const int64_t g0 = (int64_t)(t * GEN_SPAN) - GEN_LOOKBACK + 16 * tid;
    if (g0 >= (int64_t)a.first && (uint64_t)g0 + 16 <= a.end) return __ldg(reinterpret_cast<const uint4 *>(a.bases + g0));
    unsigned char c[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int64_t g = g0 + i;
      c[i] = (g >= (int64_t)a.first && (uint64_t)g < a.end) ? __ldg(a.bases + g) : (unsigned char)'\n';
    }
    uint4 v;
    memcpy(&v, c, 16);
    return v;
  };
  uint4 vnext = make_uint4(0, 0, 0, 0);
  if (blockIdx.x < a.n_tiles) vnext = fetch(blockIdx.x);

  for (uint64_t tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
    // bit arrays zero (the previous tile's readers are behind the barrier at the end of the loop body)
    for (int i = tid; i < GEN_PC_WORDS + 2 * (GEN_BITS / 32); i += GEN_THREADS) pc[i] = 0;
    if (tid == 0) { s_firstbreak = 0xffffffffu; s_base = a.tile_read[tile]; }
    // byte positions (relative to the aligned origin) [lo, hi): lo is negative for tile 0
    const int64_t lo = (int64_t)(tile * GEN_SPAN) - GEN_LOOKBACK;
    const uint4 v = vnext;
    if (tile + gridDim.x < a.n_tiles) vnext = fetch(tile + gridDim.x);          // in flight while this tile is processed

    // ---- 1. classify: four bytes at a time --------------------------------------------------------------------------------
    uint32_t clean = 0, nmask = 0;
    uint64_t codes = 0;                                                         // B bits per clean base, 16 bases at most
    if constexpr (B != 2) {
      unsigned char c[16];
      memcpy(c, &v, 16);
      int m = 0;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        if (c[i] != '\n' && c[i] != '\r') {
          clean |= 1u << i;
          codes |= (uint64_t)alpha_code<B>(c[i]) << (B * m);
          nmask |= ((c[i] == 'N' || c[i] == 'n') ? 1u : 0u) << m;
          ++m;
        }
      }
    } else {
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
      uint32_t skip16 = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t u = w[q] & 0xDFDFDFDFu;                                  // case folded (only compared with letters)
        const uint32_t eC = eq4(u, 0x43434343u), eG = eq4(u, 0x47474747u), eT = eq4(u, 0x54545454u), eN = eq4(u, 0x4E4E4E4Eu);
        const uint32_t eol = eq4(w[q], 0x0A0A0A0Au) | eq4(w[q], 0x0D0D0D0Du);
        skip16 |= mask4(eol) << (4 * q);
        codes |= (uint64_t)((spread4(eC | eT) | (spread4(eG | eT) << 1)) << (8 * q));   // DNA::FROM_ASCII: A 0, C 1, G 2, T 3, all else 0
        nmask |= mask4(eN) << (4 * q);
      }
      clean = ~skip16 & 0xffffu;
      if (skip16) {                                                             // squeeze the EOL positions out
        uint32_t cc = 0, nn = 0;
        int m = 0;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          if ((clean >> i) & 1u) {
            cc |= ((uint32_t)(codes >> (2 * i)) & 3u) << (2 * m);
            nn |= ((nmask >> i) & 1u) << m;
            ++m;
          }
        }
        codes = cc; nmask = nn;
      }
    }
    const uint32_t nclean = __popc(clean);
    uint32_t cb, ctotal;
    if (__syncthreads_or(clean != 0xffffu)) Scan(scan_tmp).ExclusiveSum(nclean, cb, ctotal);
    else { cb = 16u * (uint32_t)tid; ctotal = (uint32_t)GEN_TILE_BYTES; }          // no EOL character in the tile: positions are byte offsets
    cbase[tid] = cb; cmask[tid] = clean;
    if (nclean) {
      const uint32_t bit = (uint32_t)B * (cb + GEN_PAD);
      const uint32_t sh0 = bit & 31u;
      const uint64_t vlo = codes << sh0;                                        // 64 code bits shifted by < 32: up to three words
      const uint32_t vhi = sh0 ? (uint32_t)(codes >> (64 - sh0)) : 0u;
      atomicOr(&pc[bit >> 5], (uint32_t)vlo);
      if (vlo >> 32) atomicOr(&pc[(bit >> 5) + 1], (uint32_t)(vlo >> 32));
      if (vhi) atomicOr(&pc[(bit >> 5) + 2], vhi);
      if (nmask) {
        const uint32_t nbit = cb + GEN_PAD;
        const uint64_t nv = (uint64_t)nmask << (nbit & 31u);
        atomicOr(&nb[nbit >> 5], (uint32_t)nv);
        if (nv >> 32) atomicOr(&nb[(nbit >> 5) + 1], (uint32_t)(nv >> 32));
        atomicMin(&s_firstbreak, cb + (uint32_t)(__ffs(nmask) - 1));
      }
    }
    __syncthreads();

    // ---- 2. read starts inside the tile -> start bits in compact space -------------------------------------------------------
    {
      const uint64_t hi = (uint64_t)(lo + GEN_TILE_BYTES);
      for (uint64_t r = s_base + tid; r <= a.n_reads; r += GEN_THREADS) {
        const uint64_t b = a.offsets[r] + a.delta;
        if (b >= hi) break;                                   // offsets ascend
        const uint32_t local = (uint32_t)((int64_t)b - lo);   // 0 <= local < GEN_TILE_BYTES
        const uint32_t th = local >> 4, within = local & 15u;
        const uint32_t ci = cbase[th] + __popc(cmask[th] & ((1u << within) - 1u));   // clean bases in front of byte `local`
        atomicOr(&sb[(ci + GEN_PAD) >> 5], 1u << ((ci + GEN_PAD) & 31u));
        if (r < a.n_reads) atomicMin(&s_firstbreak, ci);
      }
    }
    __syncthreads();
    // compact index of the first output position (byte lo + GEN_LOOKBACK); tile 0 emits from its first byte
    const uint32_t c_lo = cbase[GEN_LOOKBACK / 16];
    // context check: a window of an output position may reach behind compact position 0 only if data lies there
    if (tid == 0 && lo > (int64_t)a.first && c_lo < (uint32_t)(k - 1) && s_firstbreak > c_lo && ctotal > c_lo) *a.fallback = 1;

    // ---- 3. validity of this thread's GEN_RUN windows ---------------------------------------------------------------------------
    const uint32_t j0 = c_lo + (uint32_t)tid * GEN_RUN;       // first compact position of the run
    uint32_t vmask = 0;
    uint32_t len0 = 0;                                        // good run length at j0 - 1
    const bool active = j0 < ctotal;
    uint32_t nrun = 0, srun = 0;
    if (active) {
      // previous 64 positions: the run of good bases ending at j0 - 1 starts behind the last N / at the last read start;
      // position 0 of the tile counts as a start (nothing in front of it is known)
      const uint64_t pn = bits64(nb, j0 + GEN_PAD - 64), ps = bits64(sb, j0 + GEN_PAD - 64);
      const uint32_t back_n = pn ? (uint32_t)__clzll((long long)pn) : 64u;                 // good positions behind the last N
      const uint32_t back_s = ps ? (uint32_t)__clzll((long long)ps) + 1u : 65u;            // positions from the last start on
      len0 = min(min(back_n, back_s), j0);                                                 // never longer than what the tile holds
      nrun = (uint32_t)bits64(nb, j0 + GEN_PAD) & 0xffffu;
      srun = (uint32_t)bits64(sb, j0 + GEN_PAD) & 0xffffu;
      const uint32_t left = min((uint32_t)GEN_RUN, ctotal - j0);
      // segment by segment between the breaks of the run: inside a segment that starts at position s0 with L good bases in
      // front of it, window i is valid from i >= s0 + k - 1 - L on
      uint32_t breaks = nrun | srun, s0 = 0, L = len0;
      for (;;) {
        const uint32_t b = breaks ? (uint32_t)__ffs(breaks) - 1u : (uint32_t)GEN_RUN;
        const uint32_t need = (uint32_t)k - 1u > L ? (uint32_t)k - 1u - L : 0u;
        const uint32_t fv = s0 + need;
        if (fv < b) vmask |= (0xffffu >> (GEN_RUN - b)) & ~((1u << fv) - 1u);
        if (b >= (uint32_t)GEN_RUN) break;
        L = 0;
        s0 = ((nrun >> b) & 1u) ? b + 1u : b;                 // N: the run restarts behind it; read start: at it
        breaks &= breaks - 1u;
      }
      vmask &= 0xffffu >> (GEN_RUN - left);
    }
    uint32_t obase, ototal;
    Scan(scan_tmp).ExclusiveSum((uint32_t)__popc(vmask), obase, ototal);

    // ---- look-back: where this tile's output starts -------------------------------------------------------------------------------
    if (!SCATTER && tid >= GEN_THREADS - 32) {                // the last warp generates nothing (GEN_SPAN / GEN_RUN = 480 runs)
      unsigned long long sum = 0;
      if (tile == 0) {
        if (lane == 0) atomicExch(&a.state[0], GEN_ST_INC | (unsigned long long)ototal);
      } else {
        if (lane == 0) atomicExch(&a.state[tile], GEN_ST_AGG | (unsigned long long)ototal);
        int64_t i = (int64_t)tile - 1;
        for (;;) {
          const int64_t idx = i - lane;
          unsigned long long v = GEN_ST_INC;                  // in front of tile 0: inclusive prefix 0
          if (idx >= 0) {
            const volatile unsigned long long *p = a.state + idx;
            do { v = *p; } while ((v & GEN_ST_MASK) == 0);
          }
          const uint32_t inc = __ballot_sync(FULL_MASK, (v & GEN_ST_MASK) == GEN_ST_INC);
          const int firstinc = inc ? __ffs(inc) - 1 : 32;
          unsigned long long part = lane <= firstinc ? (v & ~GEN_ST_MASK) : 0ULL;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(FULL_MASK, part, o);
          sum += part;
          if (inc) break;
          i -= 32;
        }
        if (lane == 0) atomicExch(&a.state[tile], GEN_ST_INC | (sum + (unsigned long long)ototal));
      }
      if (lane == 0) s_base = sum;
    }

    // ---- 4. rolling generation into the shared-memory out tile ------------------------------------------------------------------
    if (SCATTER && tid < WIN_MAX_DST) s_cnt[tid] = 0;
    if ((SCATTER || a.out_keys) && vmask) {
      uint64_t fwd[W], rc[W];
      if constexpr (B == 2) {
        // window ending at j0 - 1, little-endian (oldest base lowest): the top k of the 32 W positions in front of j0
        uint64_t x[W];
        const uint32_t hb = 2u * (j0 + GEN_PAD - 32u * W);
        if (W == 1) {
          x[0] = bits64(pc, hb) >> (64 - 2 * k);                // k >= 1
        } else {
          const uint64_t h0 = bits64(pc, hb), h1 = bits64(pc, hb + 64);
          const int s = 128 - 2 * k;                            // 0 <= s < 64 for 33 <= k <= 64
          x[0] = s ? ((h0 >> s) | (h1 << (64 - s))) : h0;
          x[W - 1] = h1 >> s;
        }
        // forward k-mer = 2-bit groups of x reversed (newest base lowest); reverse complement = ~x
        const int sh = 64 * W - 2 * k;
        if (W == 1) fwd[0] = rev2bit64(x[0] << sh);
        else {
          // reverse the 128-bit value (x << sh): words swap and reverse
          const uint64_t y0 = sh ? (x[0] << sh) : x[0];
          const uint64_t y1 = sh ? ((x[W - 1] << sh) | (x[0] >> (64 - sh))) : x[W - 1];
          fwd[0] = rev2bit64(y1); fwd[W - 1] = rev2bit64(y0);
        }
#pragma unroll
        for (int q = 0; q < W; ++q) rc[q] = ~x[q];
        rc[W - 1] &= topmask;
      } else {
#pragma unroll
        for (int q = 0; q < W; ++q) { fwd[q] = 0; rc[q] = 0; }
      }
      // one step of both windows: the new base enters fwd at the bottom, its complement enters rc at the top
      const int tp = B * (k - 1);                               // bit position of the oldest base of fwd = newest of rc
      auto roll = [&](uint64_t c) {
        const uint64_t cc = (uint64_t)alpha_comp<B>((uint32_t)c);
        if (W == 1) {
          fwd[0] = ((fwd[0] << B) | c) & topmask;
          rc[0] = (rc[0] >> B) | (cc << tp);
        } else {
          fwd[W - 1] = ((fwd[W - 1] << B) | (fwd[0] >> (64 - B))) & topmask;
          fwd[0] = (fwd[0] << B) | c;
          rc[0] = (rc[0] >> B) | (rc[W - 1] << (64 - B));
          rc[W - 1] >>= B;
          if (tp >= 64) rc[W - 1] |= cc << (tp - 64);
          else { rc[0] |= cc << tp; if (tp + B > 64) rc[W - 1] |= cc >> (64 - tp); }
        }
      };
      if constexpr (B != 2) {                                   // start windows rolled up from the k - 1 bases in front of j0
        for (int i = k - 1; i >= 1; --i) {
          const uint32_t bit = (uint32_t)B * (j0 + GEN_PAD - (uint32_t)i);
          roll((uint64_t)(__funnelshift_r(pc[bit >> 5], pc[(bit >> 5) + 1], bit & 31u) & ((1u << B) - 1u)));
        }
      }
      const uint64_t nxt = bits64(pc, (uint32_t)B * (j0 + GEN_PAD));
      uint32_t o = obase;
#pragma unroll
      for (int i = 0; i < GEN_RUN; ++i) {
        roll((nxt >> (B * i)) & ((1u << B) - 1u));
        if ((vmask >> i) & 1u) {
          const bool use_rc = a.canonical && !kmer_less<W>(fwd, rc);
#pragma unroll
          for (int q = 0; q < W; ++q) otile[(size_t)gen_slot(o) * W + q] = use_rc ? rc[q] : fwd[q];
          ++o;
        }
      }
    }
    __syncthreads();

    // ---- SCATTER: the tile leaves for its owners -----------------------------------------------------------------------------------------
    if constexpr (SCATTER) {
      constexpr int PER = (GEN_SPAN + GEN_THREADS - 1) / GEN_THREADS;            // k-mers per thread, strided
      const uint32_t p = a.dist.p;
      const uint32_t lt_mask = (1u << lane) - 1u;
      uint32_t own_loc[PER];                                                     // owner (4 bits) | rank inside the owner's run of this tile
#pragma unroll
      for (int j = 0; j < PER; ++j) {
        const uint32_t i = (uint32_t)j * GEN_THREADS + tid;
        uint32_t own = 0xffu;
        if (i < ototal) {
          uint64_t key[W];
#pragma unroll
          for (int q = 0; q < W; ++q) key[q] = otile[(size_t)gen_slot(i) * W + q];
          own = bin_of<W>(a.dist, key);
        }
        // one shared-memory atomic per (warp, owner)
        const uint32_t peers = __match_any_sync(FULL_MASK, own);
        uint32_t start = 0;
        if (own != 0xffu && (peers & lt_mask) == 0) start = atomicAdd(&s_cnt[own], (uint32_t)__popc(peers));
        start = __shfl_sync(FULL_MASK, start, __ffs(peers) - 1);
        own_loc[j] = (own << 16) | (start + __popc(peers & lt_mask));
      }
      __syncthreads();
      if (tid < 32) {                                     // lane r < p: owner r's run -- its place in the slab and its VIRTUAL start:
        constexpr uint32_t LK = 128 / (8 * W);           // congruent to the destination address modulo a 128-byte line, so that a warp
        const uint32_t c = tid < (int)p ? s_cnt[tid] : 0u;   // of the copy-out stores whole aligned lines of the peer buffer
        const unsigned long long gb = c ? atomicAdd(&a.cursors[tid], (unsigned long long)c) : 0ULL;
        uint32_t st = 0, vs = 0;
        for (uint32_t r = 0; r < p; ++r) {
          const uint32_t cr = __shfl_sync(FULL_MASK, c, r);
          const unsigned long long gr = __shfl_sync(FULL_MASK, gb, r);
          const uint32_t ar = (uint32_t)(reinterpret_cast<uintptr_t>(a.dst.base[r]) / (8 * W) + gr) & (LK - 1);
          vs += (ar - vs) & (LK - 1);
          if (tid == (int)r) { s_lo[r] = st; s_vlo[r] = vs; s_gbase[r] = gb; }
          st += cr; vs += cr;
        }
        if (tid == 0) { s_lo[p] = st; s_vlo[WIN_MAX_DST] = vs; }
      }
      __syncthreads();
#pragma unroll
      for (int j = 0; j < PER; ++j) {
        const uint32_t i = (uint32_t)j * GEN_THREADS + tid;
        if (i < ototal) sidx[s_lo[own_loc[j] >> 16] + (own_loc[j] & 0xffffu)] = (uint16_t)i;
      }
      __syncthreads();
      // sorted tile out: consecutive threads = consecutive positions inside an owner's run
      const uint32_t v_end = s_vlo[WIN_MAX_DST];
      for (uint32_t v = tid; v < v_end; v += GEN_THREADS) {
        uint32_t r = 0;
#pragma unroll
        for (int q = 1; q < WIN_MAX_DST; ++q) r += (q < (int)p && v >= s_vlo[q]) ? 1u : 0u;
        const uint32_t off = v - s_vlo[r];
        if (off >= s_cnt[r]) continue;                   // the gap in front of the next run
        const unsigned long long gpos = s_gbase[r] + off;
        if (gpos < a.slab_cap) {
          const uint32_t src = sidx[s_lo[r] + off];
          uint64_t *o = a.dst.base[r] + gpos * W;
#pragma unroll
          for (int q = 0; q < W; ++q) o[q] = otile[(size_t)gen_slot(src) * W + q];
        }
      }
    }
    // ---- the tile leaves: coalesced stores ------------------------------------------------------------------------------------------
    if (!SCATTER && a.out_keys) {
      const uint64_t gbase = s_base;
      uint64_t room = a.out_cap > gbase ? a.out_cap - gbase : 0;
      const uint32_t cnt = (uint32_t)min((uint64_t)ototal, room);
      uint64_t *dst = a.out_keys + gbase * W;
      if (W == 2) {                                             // one 16-byte store per k-mer
        for (uint32_t i = tid; i < cnt; i += GEN_THREADS)
          *reinterpret_cast<ulonglong2 *>(dst + 2 * (size_t)i) = *reinterpret_cast<const ulonglong2 *>(otile + 2 * (size_t)gen_slot(i));
      } else {                                                  // 16-byte stores of two k-mers behind an 8-byte aligned head
        const uint32_t head = min(cnt, (uint32_t)((reinterpret_cast<uintptr_t>(dst) >> 3) & 1u));
        if (tid < (int)head) dst[tid] = otile[gen_slot(tid)];
        const uint32_t body = (cnt - head) >> 1;
        ulonglong2 *d2 = reinterpret_cast<ulonglong2 *>(dst + head);
        for (uint32_t i = tid; i < body; i += GEN_THREADS)
          d2[i] = make_ulonglong2(otile[gen_slot(head + 2 * i)], otile[gen_slot(head + 2 * i + 1)]);
        const uint32_t done = head + 2 * body;
        if (tid == 0 && done < cnt) dst[done] = otile[gen_slot(done)];
      }
    }
    __syncthreads();
  }
}

}  // namespace kh
