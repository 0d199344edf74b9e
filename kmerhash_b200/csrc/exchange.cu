// exchange.cu -- sender / receiver halves of the multi-GPU insert for one rank (one process per GPU).
//
// Replaces, for one rank of the reference's distributed map:
//   assign_count_estimate_permute   KH/distributed_batched_robinhood_map.hpp:398-625
//       rank id per key + per-rank counts + PER-DESTINATION HyperLogLog registers + permute
//   insert_p (the receiving half)   :910-1179   merge of the estimates, reserve, local insert
// The reference already ships the HLL registers with the counts so the receiver can presize its table
// without a pass over the received keys (:1024-1060).  Here the sender's single metadata pass also
// histograms every destination's keys by TABLE REGION (top bits of the spread store hash), so the
// receiver's L2-windowed insert (table.cu) needs no pass of its own before its region sort either:
//   sender   khx_plan_scatter_dev : meta pass (hash x2, 8 B/key read)  ->  rank scatter (16 B/key)
//   (keys and the p x KHX_META_WORDS metadata travel by all-to-all)
//   receiver khx_insert_recv_dev  : merge registers (max) + histograms (sum) -> estimate -> reserve ->
//                                    region sort -> insert
#include <algorithm>
#include <cstring>
#include <vector>
#include "kh_common.h"
#include "kh_table_host.h"
#include "kh_hll.cuh"
#include "kh_window.cuh"
#include "kh_bulk.cuh"
#include "kh_kmergen.cuh"

namespace kh {

// metadata a sender produces per destination rank, as u64 words (so it travels as an int64 tensor):
//   [0] key count   [1 .. 1+REGION_FINE) region histogram   [1+REGION_FINE .. +HLL_M) HLL registers
constexpr int META_HIST = 1;
constexpr int META_REGS = 1 + REGION_FINE;
constexpr int META_WORDS = 1 + REGION_FINE + HLL_M;
constexpr int PLAN_MAX_P = 8;    // shared-memory histogram: p x 1024 x u32 = 32 KB

template <int W>
__global__ void __launch_bounds__(512) plan_kernel(const uint64_t *keys, uint64_t n, BinSpec dist, int store_kind,
                                                   HashParams store_hp, unsigned long long *meta) {
  extern __shared__ uint32_t sh[];                  // [p][REGION_FINE]
  const uint32_t p = dist.p;
  for (uint32_t i = threadIdx.x; i < p * REGION_FINE; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const bool h32 = hash_bits(store_kind) == 32;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t k[W];
    load_key<W>(keys, i, k);
    const uint32_t r = bin_of<W>(dist, k);
    const uint64_t h = hash_runtime<W>(store_kind, k, store_hp);
    uint32_t idx, rank;
    if (h32) {
      const uint32_t x = (uint32_t)h;
      idx = x >> (32 - HLL_P);
      rank = __clz((int)((x << HLL_P) | 0xFFFu)) + 1;
    } else {
      idx = (uint32_t)(h >> (64 - HLL_P));
      rank = __clzll((long long)((h << HLL_P) | 0xFFFULL)) + 1;
    }
    // registers live in global memory (p x 16 KB of u64 would not fit next to the histograms); the plain
    // load may be stale, which only costs a redundant atomicMax
    unsigned long long *reg = meta + (uint64_t)r * META_WORDS + META_REGS + idx;
    if (*reg < rank) atomicMax(reg, (unsigned long long)rank);
    atomicAdd(&sh[r * REGION_FINE + (uint32_t)(spread_hash(h) >> (64 - REGION_FINE_BITS))], 1u);
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < p * REGION_FINE; i += blockDim.x)
    if (sh[i]) atomicAdd(&meta[(uint64_t)(i / REGION_FINE) * META_WORDS + META_HIST + (i % REGION_FINE)], (unsigned long long)sh[i]);
  // keys per destination = row sums of the histogram (one warp per row)
  for (uint32_t r = threadIdx.x >> 5; r < p; r += blockDim.x >> 5) {
    uint32_t c = 0;
    for (uint32_t j = threadIdx.x & 31; j < (uint32_t)REGION_FINE; j += 32) c += sh[r * REGION_FINE + j];
    c = warp_sum(c);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&meta[(uint64_t)r * META_WORDS], (unsigned long long)c);
  }
}

// cursors[r] = keys of ranks < r, from the metadata counts
__global__ void plan_cursor_kernel(const unsigned long long *meta, uint32_t p, unsigned long long *cursors) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    unsigned long long acc = 0;
    for (uint32_t r = 0; r < p; ++r) { cursors[r] = acc; acc += meta[(uint64_t)r * META_WORDS]; }
  }
}

// receiver: fine[b] = sum over sources, regs[i] = max over sources, total = sum of counts
__global__ void merge_meta_kernel(const unsigned long long *recv_meta, uint32_t p, unsigned long long *fine,
                                  uint32_t *regs, unsigned long long *total) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (uint32_t)REGION_FINE) {
    unsigned long long s = 0;
    for (uint32_t g = 0; g < p; ++g) s += recv_meta[(uint64_t)g * META_WORDS + META_HIST + i];
    fine[i] = s;
  }
  if (i < (uint32_t)HLL_M) {
    unsigned long long m = 0;
    for (uint32_t g = 0; g < p; ++g) m = max(m, recv_meta[(uint64_t)g * META_WORDS + META_REGS + i]);
    regs[i] = (uint32_t)m;
  }
  if (i == 0) {
    unsigned long long s = 0;
    for (uint32_t g = 0; g < p; ++g) s += recv_meta[(uint64_t)g * META_WORDS];
    *total = s;
  }
}

// ---- single-pass sender: metadata + rank scatter into fixed per-sender SLABS of the owners' receive buffers -------
// khx_plan_dev + khx_scatter_peers_dev read and dist-hash every key twice because the exact receive offsets need every
// sender's counts first.  With a fixed slab per (owner, sender) pair -- sender g writes its keys for owner r at
// recv_r[g * slab_cap ...] -- no offset depends on another rank, so ONE pass does it all: dist hash -> owner, store hash ->
// HLL registers + region histogram of that owner (shared memory), tile sorted by owner in shared memory, runs stored to
// the owners over NVLink.  A slab that would overflow is cut off (the true count still goes into the metadata, so every
// rank sees it after the metadata all-gather and all ranks redo the batch through the exact two-pass path).
constexpr int PS_BLOCK = 512;
constexpr int PS_TILE_BYTES = 16384;
// shared memory: sorted tile | owner of every sorted position | META: [p][1024] histogram + [p][4096] register bytes |
// PERM: position of every sorted key inside the unsorted tile
// The sorted tile is laid out with every owner's run starting at the SAME offset inside a 128-byte line as its destination in
// the owner's slab (gaps of < 128 bytes between the runs), so that a warp's copy-out store covers whole aligned lines of
// the peer buffer: measured with tools/peer_write_bw.cu, stores that straddle lines reach 546 GB/s over NVLink, aligned ones
// 711 GB/s.  ps_vtile = positions of the padded tile.
template <int W> constexpr int ps_line_keys() { return 128 / (8 * W); }
template <int W> constexpr int ps_vtile() { return PS_TILE_BYTES / (8 * W) + WIN_MAX_DST * ps_line_keys<W>(); }
template <int W> constexpr size_t ps_smem_bytes(uint32_t p, bool meta, bool perm) {
  return (size_t)ps_vtile<W>() * 8 * W + ps_vtile<W>() + (meta ? (size_t)p * REGION_FINE * 4 + (size_t)p * HLL_M : 0) +
         (perm ? (size_t)2 * ps_vtile<W>() : 0);
}
// META: the insert sender (metadata per owner into `meta`).  PERM: the query sender -- out_perm[r * slab_cap + i] = index
// (in `keys`) of the i-th key stored into owner r's slab, so the answers, which come back in slab order, can be put
// into input order.  cursors[r] ends as the number of keys this rank has for owner r (also beyond slab_cap).
// ALIGN: pad the runs of the sorted tile to the line offsets of their destinations (see ps_vtile; pays over NVLink, not for a
// destination in this GPU's own memory)
template <int W, bool META, bool PERM, bool ALIGN = true>
__global__ void __launch_bounds__(PS_BLOCK, 3) scatter_slabs_kernel(const uint64_t *keys, uint64_t n, BinSpec dist, int store_kind,
                                                                     HashParams store_hp, const __grid_constant__ BinDst dst,
                                                                     unsigned long long slab_cap, unsigned long long *cursors,
                                                                     unsigned long long *meta, uint32_t *out_perm) {
  constexpr int TILE = PS_TILE_BYTES / (8 * W);
  constexpr int KPT = TILE / PS_BLOCK;               // 4 (W=1) / 2 (W=2)
  constexpr int VT = ps_vtile<W>(), LK = ps_line_keys<W>();
  static_assert(VT % 4 == 0, "owner bytes are cleared a word at a time");
  extern __shared__ __align__(16) unsigned char ps_smem[];
  uint64_t *skeys = reinterpret_cast<uint64_t *>(ps_smem);
  uint8_t *sown = ps_smem + (size_t)VT * 8 * W;
  uint32_t *hist = reinterpret_cast<uint32_t *>(sown + VT);                // [p][REGION_FINE]
  uint32_t *regs = hist + (META ? dist.p * REGION_FINE : 0);               // [p][HLL_M] bytes, four to a word
  uint16_t *ssrc = reinterpret_cast<uint16_t *>(regs + (META ? dist.p * (HLL_M / 4) : 0));
  __shared__ uint32_t cnt[WIN_MAX_DST], lo[WIN_MAX_DST], vend;
  __shared__ unsigned long long gbase[WIN_MAX_DST];
  const uint32_t p = dist.p;
  const int lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  if constexpr (META)
    for (uint32_t i = threadIdx.x; i < p * REGION_FINE + p * (HLL_M / 4); i += PS_BLOCK) hist[i] = 0;
  const bool h32 = hash_bits(store_kind) == 32;
  const uint64_t n_tiles = (n + TILE - 1) / TILE;
  unsigned long long my_total = 0;                   // thread r < p: keys of this CTA for owner r

  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const uint64_t base = tile * TILE;
    const int tile_n = (int)min((uint64_t)TILE, n - base);
    const uint64_t next = tile + gridDim.x;
    if (next < n_tiles && threadIdx.x < PS_TILE_BYTES / 128) {            // next tile -> L2 while this one is sorted
      const uint64_t e = next * TILE + (uint64_t)threadIdx.x * (128 / (8 * W));
      if (e < n) asm volatile("prefetch.global.L2 [%0];" ::"l"(keys + e * W));
    }
    if (threadIdx.x < WIN_MAX_DST) cnt[threadIdx.x] = 0;
    if (ALIGN)
      for (int i = threadIdx.x; i < VT / 4; i += PS_BLOCK) reinterpret_cast<uint32_t *>(sown)[i] = 0xffffffffu;   // gaps: no owner
    __syncthreads();

    uint64_t k[KPT][W];
    uint32_t own[KPT], loc[KPT];
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
      const int e = j * PS_BLOCK + threadIdx.x;
      if (e < tile_n) load_key<W>(keys, base + e, k[j]);
    }
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
      const int e = j * PS_BLOCK + threadIdx.x;
      own[j] = 0xffu;
      if (e < tile_n) {
        own[j] = bin_of<W>(dist, k[j]);
        if constexpr (META) {
          const uint64_t h = hash_runtime<W>(store_kind, k[j], store_hp);
          uint32_t idx, rank;
          if (h32) {
            const uint32_t x = (uint32_t)h;
            idx = x >> (32 - HLL_P);
            rank = __clz((int)((x << HLL_P) | 0xFFFu)) + 1;
          } else {
            idx = (uint32_t)(h >> (64 - HLL_P));
            rank = __clzll((long long)((h << HLL_P) | 0xFFFULL)) + 1;
          }
          // byte-wide max inside a shared-memory word
          uint32_t *w = regs + (own[j] * HLL_M + idx) / 4;
          const int sh = (idx & 3) * 8;
          uint32_t old = *w;
          while (((old >> sh) & 0xffu) < rank) {
            const uint32_t seen = atomicCAS(w, old, (old & ~(0xffu << sh)) | (rank << sh));
            if (seen == old) break;
            old = seen;
          }
          atomicAdd(&hist[own[j] * REGION_FINE + (uint32_t)(spread_hash(h) >> (64 - REGION_FINE_BITS))], 1u);
        }
      }
      // position inside the owner's run of this tile: one shared-memory atomic per (warp, owner)
      const uint32_t peers = __match_any_sync(FULL_MASK, own[j]);
      uint32_t start = 0;
      if (own[j] != 0xffu && (peers & lt_mask) == 0) start = atomicAdd(&cnt[own[j]], (uint32_t)__popc(peers));
      start = __shfl_sync(FULL_MASK, start, __ffs(peers) - 1);
      loc[j] = start + __popc(peers & lt_mask);
    }
    __syncthreads();

    unsigned long long gb_pending = 0;                // !ALIGN: warp 0's slab positions, not yet consumed
    if (threadIdx.x < 32) {                          // lane r < p: owner r's run -- its place in the slab, then in the (padded) tile
      const uint32_t c = threadIdx.x < p ? cnt[threadIdx.x] : 0u;
      const unsigned long long gb = c ? atomicAdd(&cursors[threadIdx.x], (unsigned long long)c) : 0ULL;
      my_total += c;
      uint32_t st = 0;
      if (ALIGN) {                                    // run r starts at the first position >= the end of run r-1 that is congruent
        for (uint32_t r = 0; r < p; ++r) {            // to its destination address modulo a line
          const uint32_t cr = __shfl_sync(FULL_MASK, c, r);
          const unsigned long long gr = __shfl_sync(FULL_MASK, gb, r);
          const uint32_t ar = (uint32_t)(reinterpret_cast<uintptr_t>(dst.base[r]) / (8 * W) + gr) & (uint32_t)(LK - 1);
          st += (ar - st) & (uint32_t)(LK - 1);
          if (threadIdx.x == r) { lo[r] = st; gbase[r] = gb; }
          st += cr;
        }
      } else {
        // unpadded: the runs' places in the tile do not depend on the slab positions, so the atomics' round trip is left pending
        // across the barrier and hides behind the shared-memory scatter of all warps; warp 0 stores the positions afterwards
        for (uint32_t r = 0; r < p; ++r) {
          const uint32_t cr = __shfl_sync(FULL_MASK, c, r);
          if (threadIdx.x == r) lo[r] = st;
          st += cr;
        }
        gb_pending = gb;
      }
      if (threadIdx.x == 0) vend = st;
    }
    __syncthreads();
    if (!ALIGN && threadIdx.x < p) gbase[threadIdx.x] = gb_pending;      // read by the copy-out, behind the next barrier

#pragma unroll
    for (int j = 0; j < KPT; ++j) {
      const int e = j * PS_BLOCK + threadIdx.x;
      if (e < tile_n) {
        const uint32_t pos = lo[own[j]] + loc[j];
#pragma unroll
        for (int q = 0; q < W; ++q) skeys[pos * W + q] = k[j][q];
        sown[pos] = (uint8_t)own[j];
        if constexpr (PERM) ssrc[pos] = (uint16_t)e;
      }
    }
    __syncthreads();

    // sorted tile out: consecutive threads = consecutive addresses inside an owner's run
    const int v_end = (int)vend;                    // a warp = whole lines of the slab
    for (int e = threadIdx.x; e < v_end; e += PS_BLOCK) {
      const uint32_t r = sown[e];
      if (r == 0xffu) continue;
      const unsigned long long gpos = gbase[r] + (uint32_t)(e - lo[r]);
      if (gpos < slab_cap) {
        uint64_t *o = dst.base[r] + gpos * W;
#pragma unroll
        for (int q = 0; q < W; ++q) o[q] = skeys[e * W + q];
        if constexpr (PERM) out_perm[(unsigned long long)r * slab_cap + gpos] = (uint32_t)(base + ssrc[e]);
      }
    }
    __syncthreads();
  }

  if constexpr (META) {
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < p * REGION_FINE; i += PS_BLOCK)
      if (hist[i]) atomicAdd(&meta[(uint64_t)(i / REGION_FINE) * META_WORDS + META_HIST + (i % REGION_FINE)], (unsigned long long)hist[i]);
    for (uint32_t i = threadIdx.x; i < p * HLL_M; i += PS_BLOCK) {
      const uint32_t v = (regs[i / 4] >> ((i & 3) * 8)) & 0xffu;
      if (v) atomicMax(&meta[(uint64_t)(i / HLL_M) * META_WORDS + META_REGS + (i % HLL_M)], (unsigned long long)v);
    }
    if (threadIdx.x < p && my_total) atomicAdd(&meta[(uint64_t)threadIdx.x * META_WORDS], my_total);
  }
}

// ---- barrier + count exchange over peer memory ------------------------------------------------------------------------
// Every receive buffer ends in a header { flags[8]; counts[8][8] }.  One tiny kernel per rank: thread r publishes this
// rank's row of the count matrix in rank r's header, fences, raises flags[my_rank] = epoch there, then waits until rank r's
// flag in the LOCAL header has reached the epoch.  When the kernel ends, every peer's stores that preceded its flag (the
// keys / answers of the kernels before it on its stream) are visible here -- a barrier without NCCL's launch latency, and
// the p x p counts arrive with it.  A peer that never shows up trips the time-out instead of hanging the GPU.
constexpr unsigned long long PEER_POISON = ~0ULL;
struct PeerHeader { unsigned long long flags[WIN_MAX_DST]; unsigned long long counts[2][WIN_MAX_DST][WIN_MAX_DST]; };
struct PeerHeaders { PeerHeader *h[WIN_MAX_DST]; };
// KH_PEER_TIMEOUT_MS (read per call: tests shorten the 10 s)
static unsigned long long peer_timeout_ns() {
  const char *e = getenv("KH_PEER_TIMEOUT_MS");
  return (e && atoll(e) > 0) ? (unsigned long long)atoll(e) * 1000000ULL : 10000000000ULL;
}
// slot: which of the two count matrices this call writes (they alternate, so a rank that is one call ahead never
// overwrites a matrix a slower rank has not copied out yet).  cstride: distance between two entries of counts_row.
__global__ void peer_sync_kernel(const __grid_constant__ PeerHeaders hd, uint32_t rank, uint32_t p, unsigned long long epoch,
                                 const unsigned long long *counts_row, int cstride, int slot, int *timed_out,
                                 unsigned long long timeout_ns) {
  const uint32_t r = threadIdx.x;
  if (r >= p) return;
  if (counts_row)
    for (uint32_t c = 0; c < p; ++c)
      *reinterpret_cast<volatile unsigned long long *>(&hd.h[r]->counts[slot][rank][c]) = counts_row[(size_t)c * cstride];
  __threadfence_system();
  *reinterpret_cast<volatile unsigned long long *>(&hd.h[r]->flags[rank]) = epoch;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  const volatile unsigned long long *mine = &hd.h[rank]->flags[r];
  bool gave_up = false;
  for (;;) {
    const unsigned long long v = *mine;
    if (v == PEER_POISON) { *timed_out = 2; break; }              // a peer gave up: fail the same barrier everywhere
    if (v >= epoch) break;
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > timeout_ns) { *timed_out = 1; gave_up = true; break; }
    __nanosleep(200);
  }
  __threadfence_system();
  // one thread saw a time-out: every peer's copy of this rank's flag becomes the poison value, so that a peer that arrives
  // late (or waits in a later barrier) fails too instead of pairing this epoch with the wrong call
  const unsigned any = __ballot_sync((1u << p) - 1u, gave_up);     // p <= 8 threads are here
  if (any) {
    *reinterpret_cast<volatile unsigned long long *>(&hd.h[r]->flags[rank]) = PEER_POISON;
    __threadfence_system();
  }
}

// insert sender: metadata block r -> owner r's metadata area (slot of this rank), by peer stores
struct MetaDst { unsigned long long *d[WIN_MAX_DST]; };
__global__ void push_meta_kernel(const unsigned long long *meta, const __grid_constant__ MetaDst dst) {
  const unsigned long long *src = meta + (size_t)blockIdx.x * META_WORDS;
  unsigned long long *o = dst.d[blockIdx.x];
  for (int i = threadIdx.x; i < META_WORDS; i += blockDim.x) o[i] = src[i];
}

// answers of a query, as they came back in slab order, into input order: out[perm[r][i]] = reply[r][i] for i < counts[r]
template <typename T>
__global__ void unpermute_slabs_kernel(const T *reply, const uint32_t *perm, const unsigned long long *counts, uint32_t p,
                                       unsigned long long slab_cap, T *out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint32_t r = 0; r < p; ++r) {
    const unsigned long long c = min(counts[r], slab_cap);
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < c; i += stride)
      out[perm[(uint64_t)r * slab_cap + i]] = reply[(uint64_t)r * slab_cap + i];
  }
}

// Behind the flag barrier of chunk c: the counts of that chunk, as they stand in the LOCAL header, become (a) the pieces
// of the receiver's level-1 launch -- piece g = sender g's sub-slab of this chunk, clamped to its capacity -- and (b) a
// private copy of the whole p x p matrix (the header slot is overwritten two barriers later).  One thread.
__global__ void make_segs_kernel(const PeerHeader *hdr, int slot, uint32_t p, uint32_t rank, unsigned long long sub_cap,
                                 unsigned long long first_key, unsigned long long slab_cap, unsigned long long tile,
                                 BkSegs *out, unsigned long long *matrix, int *overflow) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  unsigned long long acc = 0, tiles = 0;
  out->nseg = (int)p;
  for (uint32_t g = 0; g < p; ++g) {
    unsigned long long c = hdr->counts[slot][g][rank];
    if (c > sub_cap) c = sub_cap;
    out->prefix[g] = acc;
    out->base[g] = first_key + (unsigned long long)g * slab_cap;
    out->tprefix[g] = tiles;
    acc += c;
    tiles += (c + tile - 1) / tile;
  }
  out->prefix[p] = acc;
  out->tprefix[p] = tiles;
  bool over = false;
  for (uint32_t g = 0; g < p; ++g)
    for (uint32_t r = 0; r < p; ++r) {
      const unsigned long long c = hdr->counts[slot][g][r];
      matrix[g * p + r] = c;
      over = over || c > sub_cap;
    }
  if (over) *overflow = 1;
}

}  // namespace kh

using namespace kh;

extern "C" {

int khx_meta_words(void) { return META_WORDS; }

int khx_plan_scatter_dev(kh_table *t, int dist_hash, uint32_t dist_seed, uint32_t p, const uint64_t *d_keys, uint64_t n,
                         uint64_t *d_meta, uint64_t *d_out_keys, uint64_t *counts) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !d_meta || !counts || (n && (!d_keys || !d_out_keys))) { set_error("null argument"); return KH_ERR_INVALID; }
    if (p == 0 || p > (uint32_t)PLAN_MAX_P || !hash_kind_supported(dist_hash)) {
      set_error("khx_plan_scatter_dev: 1 <= p <= %d", PLAN_MAX_P); return KH_ERR_INVALID;
    }
    if (n >= (1ULL << 32)) { set_error("batch must be < 2^32 keys"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    const int dk = fold_kind(dist_hash);
    const BinSpec dist{1, 0, p, dk, make_hash_params(dk, dist_seed, t->W)};
    unsigned long long *meta = reinterpret_cast<unsigned long long *>(d_meta);
    KH_CUDA(cudaMemsetAsync(meta, 0, (size_t)p * META_WORDS * sizeof(unsigned long long), s));
    for (uint32_t r = 0; r < p; ++r) counts[r] = 0;
    if (n) {
      const size_t smem = (size_t)p * REGION_FINE * sizeof(uint32_t);
      const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 2047) / 2048, (uint64_t)t->sms * 4));
      if (t->W == 1) KH_LAUNCH((plan_kernel<1>), grid, 512, smem, s, d_keys, n, dist, t->hash_kind, t->hp, meta);
      else KH_LAUNCH((plan_kernel<2>), grid, 512, smem, s, d_keys, n, dist, t->hash_kind, t->hp, meta);
      DevBuf<unsigned long long> cursors(p);
      KH_LAUNCH(plan_cursor_kernel, 1, 32, 0, s, meta, p, cursors.p);
      std::vector<unsigned long long> h(p);
      KH_CUDA(cudaMemcpy2DAsync(h.data(), sizeof(unsigned long long), meta, META_WORDS * sizeof(unsigned long long),
                                sizeof(unsigned long long), p, cudaMemcpyDeviceToHost, s));
      if (t->W == 1) KH_LAUNCH((window_scatter_kernel<1, false>), t->sms * 4, WIN_BLOCK, 0, s, d_keys, n, dist, cursors.p, d_out_keys, nullptr);
      else KH_LAUNCH((window_scatter_kernel<2, false>), t->sms * 4, WIN_BLOCK, 0, s, d_keys, n, dist, cursors.p, d_out_keys, nullptr);
      KH_CUDA(cudaStreamSynchronize(s));
      for (uint32_t r = 0; r < p; ++r) counts[r] = h[r];
    } else {
      KH_CUDA(cudaStreamSynchronize(s));
    }
    return KH_OK;
  });
}

// ---- receive buffers shared between the processes of one box (CUDA IPC over NVLink) --------------
struct khx_peer {
  int device = 0, W = 1;
  uint32_t rank = 0, p = 1;
  uint64_t capacity = 0;                 // keys
  uint64_t *local = nullptr;             // this rank's receive buffer (cudaMalloc: IPC cannot export pool memory)
  uint64_t count_syncs = 0;              // barriers that carried a count matrix (selects the matrix slot; same on every rank)
  uint32_t reply_bytes = 0;              // bytes per key of the answer region that follows the keys (query exchange), 0 = none
  bool poisoned = false;                 // a barrier timed out: the epochs of the ranks no longer agree, every later call fails
  bool imports_closed = false;           // khx_peer_close_imports ran (teardown step 1)
  uint64_t *remote[WIN_MAX_DST] = {nullptr};
  bool opened[WIN_MAX_DST] = {false};    // mapped with cudaIpcOpenMemHandle (must be closed)
  // copy-engine schedule (khx_insert_ce_dev): its stream and the pinned per-piece count rows, made on first use
  cudaStream_t copy_stream = nullptr;    // carries the barrier of every piece
  cudaStream_t link_stream[WIN_MAX_DST] = {nullptr};   // one per destination: the copies of a piece run on several copy engines at once
  cudaEvent_t link_done[WIN_MAX_DST] = {nullptr};
  unsigned long long *h_piece_counts = nullptr;   // [16][WIN_MAX_DST]
};

static khx_peer *peer_create(int device, uint32_t rank, uint32_t p, uint64_t capacity_keys, int key_words, uint32_t reply_bytes);
khx_peer *khx_peer_create(int device, uint32_t rank, uint32_t p, uint64_t capacity_keys, int key_words) {
  return peer_create(device, rank, p, capacity_keys, key_words, 0);
}
// receive buffer of the QUERY exchange: capacity_keys keys followed by 4 answer bytes per key, one allocation / IPC handle
khx_peer *khx_peer_create_q(int device, uint32_t rank, uint32_t p, uint64_t capacity_keys, int key_words) {
  return peer_create(device, rank, p, capacity_keys, key_words, 4);
}
static khx_peer *peer_create(int device, uint32_t rank, uint32_t p, uint64_t capacity_keys, int key_words, uint32_t reply_bytes) {
  khx_peer *x = nullptr;
  int64_t rc = guarded([&]() -> int64_t {
    if (p == 0 || p > (uint32_t)WIN_MAX_DST || rank >= p || key_words < 1 || key_words > 2) {
      set_error("khx_peer_create: 1 <= p <= %d, rank < p", WIN_MAX_DST); return KH_ERR_INVALID;
    }
    DeviceGuard g(device);
    x = new khx_peer();
    x->device = device; x->W = key_words; x->rank = rank; x->p = p; x->capacity = std::max<uint64_t>(capacity_keys, 1);
    x->reply_bytes = reply_bytes;
    const size_t body = x->capacity * (8 * key_words + reply_bytes);
    // keys | answers | header | metadata area: 2 halves x p senders x META_WORDS (insert exchange)
    KH_CUDA(cudaMalloc((void **)&x->local, body + sizeof(PeerHeader) + (size_t)2 * p * META_WORDS * 8));
    KH_CUDA(cudaMemset(reinterpret_cast<unsigned char *>(x->local) + body, 0, sizeof(PeerHeader)));
    x->remote[rank] = x->local;
    return KH_OK;
  });
  if (rc != KH_OK) { delete x; return nullptr; }
  return x;
}

// Teardown is a three-step collective (CUDA: freeing exported memory while another process still has it open is undefined):
// every rank khx_peer_close_imports -> barrier (the caller's: torch.distributed) -> every rank khx_peer_destroy.
void khx_peer_close_imports(khx_peer *x) {
  if (!x || x->imports_closed) return;
  int prev = -1; cudaGetDevice(&prev); cudaSetDevice(x->device);
  cudaDeviceSynchronize();
  for (uint32_t r = 0; r < x->p; ++r) if (x->opened[r]) { cudaIpcCloseMemHandle(x->remote[r]); x->opened[r] = false; x->remote[r] = nullptr; }
  x->imports_closed = true;
  if (prev >= 0) cudaSetDevice(prev);
}
int khx_peer_poisoned(const khx_peer *x) { return (x && x->poisoned) ? 1 : 0; }

void khx_peer_destroy(khx_peer *x) {
  if (!x) return;
  khx_peer_close_imports(x);
  int prev = -1; cudaGetDevice(&prev); cudaSetDevice(x->device);
  if (x->local) cudaFree(x->local);
  if (x->copy_stream) cudaStreamDestroy(x->copy_stream);
  for (auto &st : x->link_stream) if (st) cudaStreamDestroy(st);
  for (auto &e : x->link_done) if (e) cudaEventDestroy(e);
  if (x->h_piece_counts) cudaFreeHost(x->h_piece_counts);
  if (prev >= 0) cudaSetDevice(prev);
  delete x;
}

uint64_t khx_peer_capacity(const khx_peer *x) { return x ? x->capacity : 0; }
uint64_t *khx_peer_recv_ptr(khx_peer *x) { return x ? x->local : nullptr; }

int khx_peer_handle(khx_peer *x, void *out64) {
  return (int)guarded([&]() -> int64_t {
    if (!x || !out64) { set_error("null argument"); return KH_ERR_INVALID; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    DeviceGuard g(x->device);
    cudaIpcMemHandle_t h;
    KH_CUDA(cudaIpcGetMemHandle(&h, x->local));
    memcpy(out64, &h, 64);
    return KH_OK;
  });
}

int khx_peer_open(khx_peer *x, uint32_t r, const void *handle64) {
  return (int)guarded([&]() -> int64_t {
    if (!x || !handle64 || r >= x->p || r == x->rank) { set_error("bad peer index"); return KH_ERR_INVALID; }
    DeviceGuard g(x->device);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void *ptr = nullptr;
    KH_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    x->remote[r] = (uint64_t *)ptr; x->opened[r] = true;
    return KH_OK;
  });
}

// same-process peers (tests; one process driving several GPUs with peer access enabled): a plain device pointer
int khx_peer_set(khx_peer *x, uint32_t r, uint64_t *d_ptr) {
  if (!x || r >= x->p || !d_ptr) { set_error("bad peer index"); return KH_ERR_INVALID; }
  x->remote[r] = d_ptr; x->opened[r] = false;
  return KH_OK;
}

// metadata pass only (the first half of khx_plan_scatter_dev)
int khx_plan_dev(kh_table *t, int dist_hash, uint32_t dist_seed, uint32_t p, const uint64_t *d_keys, uint64_t n,
                 uint64_t *d_meta) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !d_meta || (n && !d_keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    if (p == 0 || p > (uint32_t)PLAN_MAX_P || !hash_kind_supported(dist_hash)) {
      set_error("khx_plan_dev: 1 <= p <= %d", PLAN_MAX_P); return KH_ERR_INVALID;
    }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    const int dk = fold_kind(dist_hash);
    const BinSpec dist{1, 0, p, dk, make_hash_params(dk, dist_seed, t->W)};
    unsigned long long *meta = reinterpret_cast<unsigned long long *>(d_meta);
    KH_CUDA(cudaMemsetAsync(meta, 0, (size_t)p * META_WORDS * sizeof(unsigned long long), s));
    if (n) {
      const size_t smem = (size_t)p * REGION_FINE * sizeof(uint32_t);
      const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 2047) / 2048, (uint64_t)t->sms * 4));
      if (t->W == 1) KH_LAUNCH((plan_kernel<1>), grid, 512, smem, s, d_keys, n, dist, t->hash_kind, t->hp, meta);
      else KH_LAUNCH((plan_kernel<2>), grid, 512, smem, s, d_keys, n, dist, t->hash_kind, t->hp, meta);
    }
    KH_CUDA(cudaStreamSynchronize(s));
    return KH_OK;
  });
}

// Fused partition + exchange: the rank scatter stores every key DIRECTLY into the receive buffer of its owner
// (peer memory over NVLink), at dst_offsets[r] keys from the start of rank r's buffer -- there is no send
// buffer and no separate all-to-all.  The caller orders the ranks: every receive buffer must be free before
// (a collective that follows the previous insert) and a barrier must follow before the owners read.
int khx_scatter_peers_dev(kh_table *t, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n,
                          khx_peer *x, const uint64_t *dst_offsets, const uint64_t *dst_counts) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !x || !dst_offsets || !dst_counts || (n && !d_keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    if (!hash_kind_supported(dist_hash) || n >= (1ULL << 32) || x->W != t->W) { set_error("bad arguments"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    const int dk = fold_kind(dist_hash);
    const BinSpec dist{1, 0, x->p, dk, make_hash_params(dk, dist_seed, t->W)};
    BinDst dst{};
    for (uint32_t r = 0; r < x->p; ++r) {
      if (!x->remote[r]) { set_error("peer %u not mapped", r); return KH_ERR_INVALID; }
      if (dst_offsets[r] + dst_counts[r] > x->capacity) { set_error("receive buffer of rank %u too small", r); return KH_ERR_NOMEM; }
      dst.base[r] = x->remote[r] + dst_offsets[r] * t->W;
    }
    if (n) {
      DevBuf<unsigned long long> cursors(x->p);
      KH_CUDA(cudaMemsetAsync(cursors.p, 0, x->p * sizeof(unsigned long long), s));
      if (t->W == 1) KH_LAUNCH((window_scatter_kernel<1, false, true>), t->sms * 4, WIN_BLOCK, 0, s, d_keys, n, dist, cursors.p, nullptr, nullptr, dst);
      else KH_LAUNCH((window_scatter_kernel<2, false, true>), t->sms * 4, WIN_BLOCK, 0, s, d_keys, n, dist, cursors.p, nullptr, nullptr, dst);
    }
    KH_CUDA(cudaStreamSynchronize(s));
    return KH_OK;
  });
}

// Single-pass sender (scatter_slabs_kernel): this rank's keys for owner r go to slab `rank` of half `parity` of
// r's receive buffer, i.e. to remote[r] + ((parity * p + rank) * slab_cap) keys; d_meta receives the p metadata blocks.
// The receive buffers hold 2 * p * slab_cap keys: two halves used alternately, so a rank may start batch i+1 while
// another still reads batch i (the metadata all-gather of batch i+1 orders batch i+2 behind it).
int khx_plan_scatter_slabs_dev(kh_table *t, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n, khx_peer *x,
                               uint64_t slab_cap, int parity, uint64_t *d_meta) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !x || !d_meta || (n && !d_keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    if (!hash_kind_supported(dist_hash) || n >= (1ULL << 32) || x->W != t->W || x->p > (uint32_t)PLAN_MAX_P || parity < 0 || parity > 1) {
      set_error("khx_plan_scatter_slabs_dev: bad arguments"); return KH_ERR_INVALID;
    }
    if (2ULL * x->p * slab_cap > x->capacity) { set_error("receive buffers hold %llu keys, 2 x %u slabs of %llu do not fit",
                                                          (unsigned long long)x->capacity, x->p, (unsigned long long)slab_cap); return KH_ERR_NOMEM; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    const uint32_t p = x->p;
    const int dk = fold_kind(dist_hash);
    const BinSpec dist{1, 0, p, dk, make_hash_params(dk, dist_seed, t->W)};
    unsigned long long *meta = reinterpret_cast<unsigned long long *>(d_meta);
    KH_CUDA(cudaMemsetAsync(meta, 0, (size_t)p * META_WORDS * sizeof(unsigned long long), s));
    if (n) {
      BinDst dst{};
      for (uint32_t r = 0; r < p; ++r) {
        if (!x->remote[r]) { set_error("peer %u not mapped", r); return KH_ERR_INVALID; }
        dst.base[r] = x->remote[r] + ((uint64_t)parity * p + x->rank) * slab_cap * t->W;
      }
      DevBuf<unsigned long long> cursors(p);
      KH_CUDA(cudaMemsetAsync(cursors.p, 0, p * sizeof(unsigned long long), s));
      const uint64_t tile = PS_TILE_BYTES / (8 * t->W);
      const size_t smem = t->W == 1 ? ps_smem_bytes<1>(p, true, false) : ps_smem_bytes<2>(p, true, false);
      const int per_sm = 3 * (smem + 1024) <= 227 * 1024 ? 3 : 2;       // resident CTAs (40 registers x 512 threads allow 3)
      const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + tile - 1) / tile, (uint64_t)t->sms * per_sm));
      if (t->W == 1) {
        KH_CUDA(cudaFuncSetAttribute((scatter_slabs_kernel<1, true, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        KH_LAUNCH((scatter_slabs_kernel<1, true, false>), grid, PS_BLOCK, smem, s, d_keys, n, dist, t->hash_kind, t->hp, dst,
                  (unsigned long long)slab_cap, cursors.p, meta, (uint32_t *)nullptr);
      } else {
        KH_CUDA(cudaFuncSetAttribute((scatter_slabs_kernel<2, true, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        KH_LAUNCH((scatter_slabs_kernel<2, true, false>), grid, PS_BLOCK, smem, s, d_keys, n, dist, t->hash_kind, t->hp, dst,
                  (unsigned long long)slab_cap, cursors.p, meta, (uint32_t *)nullptr);
      }
      KH_CUDA(cudaStreamSynchronize(s));     // cursors goes out of scope; the stores have been issued and completed
    } else {
      KH_CUDA(cudaStreamSynchronize(s));
    }
    return KH_OK;
  });
}

// The same sender without metadata (the receiver's level-1 pass computes the HyperLogLog registers of what it received
// itself, bit-identical to the merged sender registers): ONE dist hash per key, 16 B/key of traffic.  d_counts (DEVICE, p
// words) ends as the number of keys this rank has for every owner (also beyond slab_cap) = this rank's row of the count
// matrix for khx_peer_sync_dev.  Asynchronous on the table's stream; the barrier kernel that follows orders it.
int khx_scatter_slabs_dev(kh_table *t, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n, khx_peer *x,
                          uint64_t slab_cap, int parity, uint64_t *d_counts, int reset_counts) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !x || !d_counts || (n && !d_keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    if (!hash_kind_supported(dist_hash) || n >= (1ULL << 32) || x->W != t->W || x->p > (uint32_t)WIN_MAX_DST || parity < 0 || parity > 1) {
      set_error("khx_scatter_slabs_dev: bad arguments"); return KH_ERR_INVALID;
    }
    if (2ULL * x->p * slab_cap > x->capacity) { set_error("receive buffers hold %llu keys, 2 x %u slabs of %llu do not fit",
                                                          (unsigned long long)x->capacity, x->p, (unsigned long long)slab_cap); return KH_ERR_NOMEM; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    const uint32_t p = x->p;
    const int dk = fold_kind(dist_hash);
    const BinSpec dist{1, 0, p, dk, make_hash_params(dk, dist_seed, t->W)};
    unsigned long long *cursors = reinterpret_cast<unsigned long long *>(d_counts);
    if (reset_counts) KH_CUDA(cudaMemsetAsync(cursors, 0, p * sizeof(unsigned long long), s));
    if (n) {
      BinDst dst{};
      for (uint32_t r = 0; r < p; ++r) {
        if (!x->remote[r]) { set_error("peer %u not mapped", r); return KH_ERR_INVALID; }
        dst.base[r] = x->remote[r] + ((uint64_t)parity * p + x->rank) * slab_cap * t->W;
      }
      const uint64_t tile = PS_TILE_BYTES / (8 * t->W);
      const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + tile - 1) / tile, (uint64_t)t->sms * 3));
      const size_t smem = t->W == 1 ? ps_smem_bytes<1>(p, false, false) : ps_smem_bytes<2>(p, false, false);
      // KH_SENDER_UNPADDED=1: the variant the copy-engine schedule uses for its local partition (tools/sender_local_bench.py and
      // the emulated-exchange tests can reach it on one GPU this way)
      const bool unpadded = getenv("KH_SENDER_UNPADDED") && atoi(getenv("KH_SENDER_UNPADDED")) != 0;
      if (unpadded) {
        if (t->W == 1) KH_LAUNCH((scatter_slabs_kernel<1, false, false, false>), grid, PS_BLOCK, smem, s, d_keys, n, dist, t->hash_kind, t->hp, dst,
                                 (unsigned long long)slab_cap, cursors, (unsigned long long *)nullptr, (uint32_t *)nullptr);
        else KH_LAUNCH((scatter_slabs_kernel<2, false, false, false>), grid, PS_BLOCK, smem, s, d_keys, n, dist, t->hash_kind, t->hp, dst,
                       (unsigned long long)slab_cap, cursors, (unsigned long long *)nullptr, (uint32_t *)nullptr);
      } else {
        if (t->W == 1) KH_LAUNCH((scatter_slabs_kernel<1, false, false>), grid, PS_BLOCK, smem, s, d_keys, n, dist, t->hash_kind, t->hp, dst,
                                 (unsigned long long)slab_cap, cursors, (unsigned long long *)nullptr, (uint32_t *)nullptr);
        else KH_LAUNCH((scatter_slabs_kernel<2, false, false>), grid, PS_BLOCK, smem, s, d_keys, n, dist, t->hash_kind, t->hp, dst,
                       (unsigned long long)slab_cap, cursors, (unsigned long long *)nullptr, (uint32_t *)nullptr);
      }
    }
    return KH_OK;
  });
}

// k-mer generation FUSED with the sender ("fused with k-mer gen on the way in"): the tile kernel of
// kh_kmergen.cuh in SCATTER mode -- reads -> canonical k-mers in a shared-memory tile -> dist hash -> owner -> peer stores
// into this rank's slab of the owner's buffer.  The k-mers never exist as an array in this GPU's HBM (8 B written + 8 B read
// per k-mer saved).  d_counts (DEVICE, p words) as for khx_scatter_slabs_dev.  Returns KH_OK, or KH_ERR_FULL when the input
// holds more EOL characters inside a window than a tile's context (nothing usable was stored: the caller generates the
// k-mers with kh_kmers_from_reads_dev and sends them with khx_scatter_slabs_dev).  Synchronises the table's stream.
int khx_kmers_scatter_slabs_dev(kh_table *t, const char *d_bases, const uint64_t *d_offsets, uint64_t n_reads, int k, int canonical,
                                int dist_hash, uint32_t dist_seed, khx_peer *x, uint64_t slab_cap, int parity, uint64_t *d_counts) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !x || !d_counts || (n_reads && (!d_bases || !d_offsets))) { set_error("null argument"); return KH_ERR_INVALID; }
    if (k < 1 || k > 32 * t->W || kmer_words(k) != t->W || !hash_kind_supported(dist_hash) || x->W != t->W || x->p > (uint32_t)WIN_MAX_DST ||
        parity < 0 || parity > 1 || 2ULL * x->p * slab_cap > x->capacity) {
      set_error("khx_kmers_scatter_slabs_dev: bad arguments (k must match the table's key width)"); return KH_ERR_INVALID;
    }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    const uint32_t p = x->p;
    const int W = t->W;
    KH_CUDA(cudaMemsetAsync(d_counts, 0, p * sizeof(unsigned long long), s));
    if (n_reads == 0) { KH_CUDA(cudaStreamSynchronize(s)); return KH_OK; }
    uint64_t h_off[2] = {0, 0};
    KH_CUDA(cudaMemcpyAsync(&h_off[0], d_offsets, sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
    KH_CUDA(cudaMemcpyAsync(&h_off[1], d_offsets + n_reads, sizeof(uint64_t), cudaMemcpyDeviceToHost, s));
    KH_CUDA(cudaStreamSynchronize(s));
    if (h_off[1] <= h_off[0]) return KH_OK;
    GenArgs a{};
    a.delta = (uint64_t)(reinterpret_cast<uintptr_t>(d_bases) & 15u);
    a.bases = reinterpret_cast<const unsigned char *>(d_bases) - a.delta;
    a.offsets = d_offsets; a.n_reads = n_reads;
    a.first = h_off[0] + a.delta; a.end = h_off[1] + a.delta;
    a.k = k; a.canonical = canonical;
    a.n_tiles = (a.end + GEN_SPAN - 1) / GEN_SPAN;
    a.out_keys = nullptr; a.out_cap = 0; a.state = nullptr;
    const int dk = fold_kind(dist_hash);
    a.dist = BinSpec{1, 0, p, dk, make_hash_params(dk, dist_seed, W)};
    for (uint32_t r = 0; r < p; ++r) {
      if (!x->remote[r]) { set_error("peer %u not mapped", r); return KH_ERR_INVALID; }
      a.dst.base[r] = x->remote[r] + ((uint64_t)parity * p + x->rank) * slab_cap * W;
    }
    a.slab_cap = slab_cap;
    a.cursors = reinterpret_cast<unsigned long long *>(d_counts);
    DevBuf<uint64_t> tile_read(a.n_tiles);
    DevBuf<int> flag(1);
    KH_CUDA(cudaMemsetAsync(flag.p, 0, sizeof(int), s));
    KH_LAUNCH(gen_tile_reads_kernel, (int)((a.n_tiles + 255) / 256), 256, 0, s, d_offsets, n_reads, a.delta, a.n_tiles, tile_read.p);
    a.tile_read = tile_read.p; a.fallback = flag.p;
    const size_t smem = W == 1 ? gen_smem_bytes<1>(true) : gen_smem_bytes<2>(true);
    const int per_sm = std::max<int>(1, std::min<int>(2, (int)((227 * 1024) / (smem + 3 * 1024))));
    const int grid = (int)std::min<uint64_t>(a.n_tiles, (uint64_t)t->sms * per_sm);
    if (W == 1) {
      KH_CUDA(cudaFuncSetAttribute((kmer_tiles_kernel<1, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      KH_LAUNCH((kmer_tiles_kernel<1, true>), grid, GEN_THREADS, smem, s, a);
    } else {
      KH_CUDA(cudaFuncSetAttribute((kmer_tiles_kernel<2, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      KH_LAUNCH((kmer_tiles_kernel<2, true>), grid, GEN_THREADS, smem, s, a);
    }
    int h_flag = 0;
    KH_CUDA(cudaMemcpyAsync(&h_flag, flag.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    KH_CUDA(cudaStreamSynchronize(s));
    if (h_flag) { set_error("khx_kmers_scatter_slabs_dev: EOL-dense input, use the two-step path"); return KH_ERR_FULL; }
    return KH_OK;
  });
}

// Receiver of the slab exchange: counts[g] (HOST) keys lie in slab g of half `parity` of this rank's receive buffer.
int64_t khx_insert_recv_slabs_dev(kh_table *t, khx_peer *x, uint64_t slab_cap, int parity, const uint64_t *counts,
                                  const uint64_t *d_recv_meta, uint32_t default_val) {
  return guarded([&]() -> int64_t {
    if (!t || !x || !counts) { set_error("null argument"); return KH_ERR_INVALID; }
    if (x->W != t->W || parity < 0 || parity > 1 || 2ULL * x->p * slab_cap > x->capacity) { set_error("khx_insert_recv_slabs_dev: bad arguments"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    const uint32_t p = x->p;
    BkSegs segs{};
    segs.nseg = (int)p;
    uint64_t n = 0;
    for (uint32_t q = 0; q < p; ++q) {
      if (counts[q] > slab_cap) { set_error("slab %u holds %llu keys, more than its capacity", q, (unsigned long long)counts[q]); return KH_ERR_INVALID; }
      segs.prefix[q] = n;
      segs.base[q] = ((uint64_t)parity * p + q) * slab_cap;
      n += counts[q];
    }
    segs.prefix[p] = n;
    if (n == 0) return 0;
    // the receiver reads its p slabs ONCE: the bulk insert's level-1 pass computes the HyperLogLog registers of the received
    // keys itself (= the merged sender registers, bit for bit), so the metadata is not needed
    (void)d_recv_meta;
    PreEstimate pre{nullptr, &segs};
    return insert_dev_impl(t, x->local, nullptr, n, default_val, 1, &pre);
  });
}

// ---- query exchange (find / count / erase) over the slabs -----------------------------------------------------------
// replaces count_p / find_p / erase_p's exchange (KH/distributed_batched_robinhood_map.hpp:1258-1516,1619,2169) and the
// reply of ialltoallv_and_query_one_to_one (KH/incremental_mxx.hpp:4403): keys go out by peer stores (scatter_slabs_kernel
// with PERM), the owner's lookup kernel stores every answer straight into the asker's reply region, the asker puts the
// answers into input order.  Between the steps the caller needs: the p x p counts (all-gather of d_counts, which is also
// the barrier behind the key stores) and a barrier behind the answer stores.
int khx_query_scatter_slabs_dev(kh_table *t, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n, khx_peer *x,
                                uint64_t slab_cap, uint32_t *d_perm, uint64_t *d_counts) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !x || !d_counts || (n && (!d_keys || !d_perm))) { set_error("null argument"); return KH_ERR_INVALID; }
    if (!hash_kind_supported(dist_hash) || n >= (1ULL << 32) || x->W != t->W || x->p > (uint32_t)WIN_MAX_DST || x->reply_bytes == 0) {
      set_error("khx_query_scatter_slabs_dev: bad arguments (buffers from khx_peer_create_q?)"); return KH_ERR_INVALID;
    }
    if ((uint64_t)x->p * slab_cap > x->capacity) { set_error("receive buffers hold %llu keys, %u slabs of %llu do not fit",
                                                             (unsigned long long)x->capacity, x->p, (unsigned long long)slab_cap); return KH_ERR_NOMEM; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    const uint32_t p = x->p;
    const int dk = fold_kind(dist_hash);
    const BinSpec dist{1, 0, p, dk, make_hash_params(dk, dist_seed, t->W)};
    unsigned long long *cursors = reinterpret_cast<unsigned long long *>(d_counts);
    KH_CUDA(cudaMemsetAsync(cursors, 0, p * sizeof(unsigned long long), s));
    if (n) {
      BinDst dst{};
      for (uint32_t r = 0; r < p; ++r) {
        if (!x->remote[r]) { set_error("peer %u not mapped", r); return KH_ERR_INVALID; }
        dst.base[r] = x->remote[r] + (uint64_t)x->rank * slab_cap * t->W;
      }
      const uint64_t tile = PS_TILE_BYTES / (8 * t->W);
      const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + tile - 1) / tile, (uint64_t)t->sms * 3));
      const size_t smem = t->W == 1 ? ps_smem_bytes<1>(p, false, true) : ps_smem_bytes<2>(p, false, true);
      if (t->W == 1) KH_LAUNCH((scatter_slabs_kernel<1, false, true>), grid, PS_BLOCK, smem, s, d_keys, n, dist, t->hash_kind, t->hp, dst,
                               (unsigned long long)slab_cap, cursors, (unsigned long long *)nullptr, d_perm);
      else KH_LAUNCH((scatter_slabs_kernel<2, false, true>), grid, PS_BLOCK, smem, s, d_keys, n, dist, t->hash_kind, t->hp, dst,
                     (unsigned long long)slab_cap, cursors, (unsigned long long *)nullptr, d_perm);
    }
    KH_CUDA(cudaStreamSynchronize(s));
    return KH_OK;
  });
}

// owner side: counts[g] (HOST) keys of asker g lie in slab g of this rank's buffer.  op 0 = find (u32 answers, `nonexistent`
// for absent keys), 1 = count (u8 answers), 2 = erase (no answers).  Answers go to slab `rank` of asker g's reply region.
// Returns the number of hits / erased keys.
int64_t khx_query_slabs_dev(kh_table *t, khx_peer *x, int op, uint64_t slab_cap, const uint64_t *counts, uint32_t nonexistent) {
  return guarded([&]() -> int64_t {
    if (!t || !x || !counts || op < 0 || op > 2) { set_error("bad argument"); return KH_ERR_INVALID; }
    if (x->W != t->W || x->reply_bytes == 0 || (uint64_t)x->p * slab_cap > x->capacity) { set_error("khx_query_slabs_dev: bad arguments"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    const uint32_t p = x->p;
    const int W = t->W;
    for (uint32_t q = 0; q < p; ++q)
      if (counts[q] > slab_cap) { set_error("slab %u holds %llu keys, more than its capacity", q, (unsigned long long)counts[q]); return KH_ERR_INVALID; }
    int64_t total = 0;
    if (op == 2) {                       // one erase over all slabs (the shrink rule looks at the table once, as the reference does)
      uint64_t n = 0;
      for (uint32_t q = 0; q < p; ++q) n += counts[q];
      if (!n) return 0;
      DevBuf<uint64_t> flat(n * W);
      uint64_t off = 0;
      for (uint32_t q = 0; q < p; ++q) if (counts[q]) {
        KH_CUDA(cudaMemcpyAsync(flat.p + off * W, x->local + (uint64_t)q * slab_cap * W, counts[q] * 8 * W, cudaMemcpyDeviceToDevice, t->stream[0]));
        off += counts[q];
      }
      KH_CUDA(cudaStreamSynchronize(t->stream[0]));
      return kh_erase_dev(t, flat.p, n);
    }
    // one launch per sender slab, back to back (no synchronisation in between), the answers stored straight into the askers' buffers
    const size_t elem = op == 0 ? 4 : 1;
    const uint64_t *keys[WIN_MAX_DST]; void *reply[WIN_MAX_DST]; uint64_t cnt[WIN_MAX_DST];
    for (uint32_t q = 0; q < p; ++q) {
      cnt[q] = counts[q];
      if (counts[q] && !x->remote[q]) { set_error("peer %u not mapped", q); return KH_ERR_INVALID; }
      keys[q] = x->local + (uint64_t)q * slab_cap * W;
      reply[q] = counts[q] ? reinterpret_cast<unsigned char *>(x->remote[q]) + x->capacity * 8 * W + (uint64_t)x->rank * slab_cap * elem : nullptr;
    }
    total = query_pieces_dev(t, op, (int)p, keys, cnt, reply, nonexistent);
    return total;
  });
}

// barrier over peer memory (peer_sync_kernel).  epoch: 1, 2, 3, ... per buffer set, the same on every rank.  d_counts_row
// (DEVICE, p words, may be null): this rank's row of the count matrix, published to every rank; counts_matrix (HOST, p x p,
// may be null) receives the matrix [sender][owner] as it stands in the local header after the barrier.
static int peer_sync(kh_table *t, khx_peer *x, uint64_t epoch, const uint64_t *d_counts_row, int counts_stride, uint64_t *counts_matrix) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !x || epoch == 0) { set_error("bad argument"); return KH_ERR_INVALID; }
    if (x->poisoned) { set_error("khx_peer_sync_dev: an earlier barrier on these buffers timed out; recreate them"); return KH_ERR_CUDA; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    const size_t body = x->capacity * (size_t)(8 * x->W + x->reply_bytes);
    PeerHeaders hd{};
    for (uint32_t r = 0; r < x->p; ++r) {
      if (!x->remote[r]) { set_error("peer %u not mapped", r); return KH_ERR_INVALID; }
      hd.h[r] = reinterpret_cast<PeerHeader *>(reinterpret_cast<unsigned char *>(x->remote[r]) + body);
    }
    DevBuf<int> flag(1);
    KH_CUDA(cudaMemsetAsync(flag.p, 0, sizeof(int), s));
    const int slot = (int)(x->count_syncs & 1);
    if (d_counts_row) ++x->count_syncs;
    const unsigned long long timeout_ns = peer_timeout_ns();
    KH_LAUNCH(peer_sync_kernel, 1, 32, 0, s, hd, x->rank, x->p, (unsigned long long)epoch,
              reinterpret_cast<const unsigned long long *>(d_counts_row), counts_stride, slot, flag.p, timeout_ns);
    int h_flag = 0;
    PeerHeader h_hdr;
    KH_CUDA(cudaMemcpyAsync(&h_flag, flag.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    if (counts_matrix) KH_CUDA(cudaMemcpyAsync(&h_hdr, hd.h[x->rank], sizeof(PeerHeader), cudaMemcpyDeviceToHost, s));
    KH_CUDA(cudaStreamSynchronize(s));
    if (h_flag) {
      x->poisoned = true;
      set_error(h_flag == 2 ? "khx_peer_sync_dev: a peer timed out in a barrier (epoch %llu); buffers are unusable"
                            : "khx_peer_sync_dev: a peer did not reach epoch %llu in time; buffers are unusable", (unsigned long long)epoch);
      return KH_ERR_CUDA;
    }
    if (counts_matrix)
      for (uint32_t a = 0; a < x->p; ++a)
        for (uint32_t b = 0; b < x->p; ++b) counts_matrix[a * x->p + b] = h_hdr.counts[slot][a][b];
    return KH_OK;
  });
}
int khx_peer_sync_dev(kh_table *t, khx_peer *x, uint64_t epoch, const uint64_t *d_counts_row, uint64_t *counts_matrix) {
  return peer_sync(t, x, epoch, d_counts_row, 1, counts_matrix);
}

// insert exchange without NCCL: push metadata block r (of d_meta, as khx_plan_scatter_slabs_dev filled it) into owner r's
// metadata area of half `parity`, then the barrier, publishing the counts d_meta[r][0] on the way.
int khx_push_meta_sync_dev(kh_table *t, khx_peer *x, int parity, const uint64_t *d_meta, uint64_t epoch, uint64_t *counts_matrix) {
  const int64_t rc = guarded([&]() -> int64_t {
    if (!t || !x || !d_meta || parity < 0 || parity > 1) { set_error("bad argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    const size_t body = x->capacity * (size_t)(8 * x->W + x->reply_bytes) + sizeof(PeerHeader);
    MetaDst dst{};
    for (uint32_t r = 0; r < x->p; ++r) {
      if (!x->remote[r]) { set_error("peer %u not mapped", r); return KH_ERR_INVALID; }
      dst.d[r] = reinterpret_cast<unsigned long long *>(reinterpret_cast<unsigned char *>(x->remote[r]) + body) +
                 ((size_t)parity * x->p + x->rank) * META_WORDS;
    }
    KH_LAUNCH(push_meta_kernel, x->p, 256, 0, t->stream[0], reinterpret_cast<const unsigned long long *>(d_meta), dst);
    return KH_OK;
  });
  if (rc != KH_OK) return (int)rc;
  return peer_sync(t, x, epoch, d_meta, META_WORDS, counts_matrix);
}
// this rank's metadata area of half `parity`: p blocks of khx_meta_words() words, block g from sender g
uint64_t *khx_peer_meta_ptr(khx_peer *x, int parity) {
  if (!x || parity < 0 || parity > 1) return nullptr;
  const size_t body = x->capacity * (size_t)(8 * x->W + x->reply_bytes) + sizeof(PeerHeader);
  return reinterpret_cast<uint64_t *>(reinterpret_cast<unsigned char *>(x->local) + body) + (size_t)parity * x->p * META_WORDS;
}

// asker side, after the barrier behind the answer stores: d_out[perm[r][i]] = reply[r][i]  (elem_bytes 4 or 1)
int khx_unpermute_slabs_dev(kh_table *t, khx_peer *x, int elem_bytes, uint64_t slab_cap, const uint32_t *d_perm, const uint64_t *d_counts,
                            uint64_t n, void *d_out) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !x || !d_counts || (n && (!d_perm || !d_out)) || (elem_bytes != 1 && elem_bytes != 4)) { set_error("bad argument"); return KH_ERR_INVALID; }
    if (x->reply_bytes == 0 || (uint64_t)x->p * slab_cap > x->capacity) { set_error("khx_unpermute_slabs_dev: bad arguments"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    if (n) {
      const unsigned char *reply = reinterpret_cast<const unsigned char *>(x->local) + x->capacity * 8 * x->W;
      const unsigned long long *cnt = reinterpret_cast<const unsigned long long *>(d_counts);
      const int grid = t->grid_for(std::max<uint64_t>(n / x->p, 1), 256);
      if (elem_bytes == 4) KH_LAUNCH((unpermute_slabs_kernel<uint32_t>), grid, 256, 0, s, reinterpret_cast<const uint32_t *>(reply), d_perm, cnt,
                                     x->p, (unsigned long long)slab_cap, reinterpret_cast<uint32_t *>(d_out));
      else KH_LAUNCH((unpermute_slabs_kernel<uint8_t>), grid, 256, 0, s, reply, d_perm, cnt, x->p, (unsigned long long)slab_cap,
                     reinterpret_cast<uint8_t *>(d_out));
    }
    KH_CUDA(cudaStreamSynchronize(s));
    return KH_OK;
  });
}

int64_t khx_insert_recv_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, const uint64_t *d_recv_meta, uint32_t p,
                            uint32_t default_val) {
  return guarded([&]() -> int64_t {
    if (!t || !d_recv_meta || (n && !d_keys) || p == 0) { set_error("null argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    if (n == 0) return 0;
    cudaStream_t s = t->stream[0];
    DevBuf<unsigned long long> fine(REGION_FINE), total(1);
    DevBuf<uint32_t> regs(HLL_M);
    KH_LAUNCH(merge_meta_kernel, (HLL_M + 255) / 256, 256, 0, s, reinterpret_cast<const unsigned long long *>(d_recv_meta), p,
              fine.p, regs.p, total.p);
    unsigned long long h_total = 0;
    KH_CUDA(cudaMemcpyAsync(&h_total, total.p, sizeof(h_total), cudaMemcpyDeviceToHost, s));
    uint8_t batch_regs[HLL_M];
    hll_finish(t->hash_kind, regs.p, batch_regs, s);   // synchronises s
    if (h_total != n) { set_error("received %llu keys but the metadata announces %llu", (unsigned long long)n, h_total); return KH_ERR_INVALID; }
    PreEstimate pre{batch_regs, nullptr};
    return insert_dev_impl(t, d_keys, nullptr, n, default_val, 1, &pre);
  });
}

// ---- pipelined distributed insert: the exchange overlapped with the local insertion -----------------------------------------
// The reference's ialltoallv_and_modify (KH/incremental_mxx.hpp:3437-3645) posts the transfers of all pieces and runs the
// local insert on every piece as it lands.  Here the batch is cut into `chunks` pieces; for piece c, on the send stream:
//   sender kernel (dist hash -> owner -> peer stores into sub-slab c of this rank's slab)  ->  flag barrier kernel, which
//   also delivers the p x p counts of the piece  ->  make_segs_kernel (pieces of the receive side, device resident);
// and on the receive stream, behind an event: the level-1 kernel of the bulk insert over the p sub-slabs of piece c --
// launched with a fixed grid before the host knows any count.  Nothing synchronises with the host until the last piece,
// so the level-1 pass of piece c runs WHILE the senders of piece c+1 store (the sender takes one CTA per SM, the
// level-1 kernel the other).  Then: estimate -> reserve -> tile insert as for any bulk batch.
// counts_matrix (HOST, p x p): keys sender g had for owner r, summed over the pieces.  *overflowed = 1: a sub-slab was too
// small -- the same verdict on every rank, the table is untouched and the caller repeats the batch on the exact path.
// Epochs epoch0 .. epoch0 + chunks - 1 are consumed.  Returns the number of new keys.
// ce = false: the sender kernels store into the peers' slabs themselves (SM stores over NVLink).
// ce = true:  the sender kernels partition into a LOCAL staging buffer (and, for the keys this rank owns, straight into its own
//             slab); the COPY ENGINES move every (piece, owner) run to the owner's slab while the SMs go on with the next piece's
//             partition and the previous pieces' level-1 passes.  NVLink time leaves the SMs' critical path.
static int64_t insert_pipelined_impl(kh_table *t, khx_peer *x, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n,
                                     uint64_t slab_cap, int parity, int chunks, uint64_t epoch0, uint32_t default_val,
                                     uint64_t *counts_matrix, int *overflowed, bool ce) {
  return guarded([&]() -> int64_t {
    if (!t || !x || !counts_matrix || !overflowed || (n && !d_keys) || epoch0 == 0) { set_error("null argument"); return KH_ERR_INVALID; }
    const uint32_t p = x->p;
    const int W = t->W;
    if (!hash_kind_supported(dist_hash) || n >= (1ULL << 32) || x->W != W || parity < 0 || parity > 1 || chunks < 1 || chunks > 16 ||
        slab_cap % (2ULL * chunks) != 0 || 2ULL * p * slab_cap > x->capacity || p > (uint32_t)BK_MAX_SEG) {
      set_error("khx_insert_pipelined_dev / khx_insert_ce_dev: bad arguments (slab_cap must be a multiple of 2 * chunks, 2 * p * slab_cap keys must fit)");
      return KH_ERR_INVALID;
    }
    if (x->poisoned) { set_error("khx_insert_pipelined_dev: an earlier barrier on these buffers timed out; recreate them"); return KH_ERR_CUDA; }
    if (!bulk_pipe_candidate(t, (uint64_t)p * slab_cap)) { set_error("khx_insert_pipelined_dev: the table does not take the bulk route"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    cudaStream_t S = t->stream[0], R = t->stream[1];
    const uint64_t sub_cap = slab_cap / chunks;
    const int dk = fold_kind(dist_hash);
    const BinSpec dist{1, 0, p, dk, make_hash_params(dk, dist_seed, W)};
    const size_t body = x->capacity * (size_t)(8 * x->W + x->reply_bytes);
    PeerHeaders hd{};
    for (uint32_t r = 0; r < p; ++r) {
      if (!x->remote[r]) { set_error("peer %u not mapped", r); return KH_ERR_INVALID; }
      hd.h[r] = reinterpret_cast<PeerHeader *>(reinterpret_cast<unsigned char *>(x->remote[r]) + body);
    }
    const unsigned long long timeout_ns = peer_timeout_ns();
    DevBuf<unsigned long long> d_counts((size_t)chunks * p), d_matrix((size_t)chunks * p * p);
    DevBuf<BkSegs> d_segs(chunks);
    DevBuf<int> d_flags(2);                                   // [0] barrier time-out, [1] sub-slab overflow
    KH_CUDA(cudaMemsetAsync(d_counts.p, 0, (size_t)chunks * p * sizeof(unsigned long long), S));
    KH_CUDA(cudaMemsetAsync(d_flags.p, 0, 2 * sizeof(int), S));
    std::vector<cudaEvent_t> ev(chunks, nullptr);
    struct EvGuard { std::vector<cudaEvent_t> &e; ~EvGuard() { for (auto v : e) if (v) cudaEventDestroy(v); } } evg{ev};
    // KH_TRACE=1: the events carry times; the schedule of this call is printed at its end (rank 0's view of the overlap)
    const bool tr_on = getenv("KH_TRACE") && atoi(getenv("KH_TRACE")) != 0;
    const unsigned ev_flags = tr_on ? cudaEventDefault : cudaEventDisableTiming;
    for (auto &e : ev) KH_CUDA(cudaEventCreateWithFlags(&e, ev_flags));
    std::vector<cudaEvent_t> evt(tr_on ? 2 + chunks : 0, nullptr);      // [0] start (S), [1 + c] level 1 of piece c done (R), [1 + chunks] finish
    EvGuard evgt{evt};
    for (auto &e : evt) KH_CUDA(cudaEventCreate(&e));
    if (tr_on) KH_CUDA(cudaEventRecord(evt[0], S));

    BulkBatch b;
    bulk_pipe_begin(b, t, (uint64_t)p * slab_cap, R);
    const uint64_t tile_send = PS_TILE_BYTES / (8 * W);
    const uint64_t tile_l1 = l1_tile_keys(W);
    const size_t smem = W == 1 ? ps_smem_bytes<1>(p, false, false) : ps_smem_bytes<2>(p, false, false);
    // ce: staging for the runs that leave this GPU, [owner][slab_cap] keys; the copy stream; pinned count rows; one event per piece
    DevBuf<uint64_t> stage;
    cudaStream_t C = S;                                       // the stream that carries the barrier of a piece
    bool multi_link = false;
    std::vector<cudaEvent_t> evp(ce ? chunks : 0, nullptr);
    struct EvGuard2 { std::vector<cudaEvent_t> &e; ~EvGuard2() { for (auto v : e) if (v) cudaEventDestroy(v); } } evg2{evp};
    if (ce) {
      if (!x->copy_stream) KH_CUDA(cudaStreamCreateWithFlags(&x->copy_stream, cudaStreamNonBlocking));
      // KH_CE_STREAMS=1 (default): all copies of a rank on ONE stream, destinations rotated by rank, so that at any time every GPU
      // sends to one peer and receives from one.  KH_CE_STREAMS=0: one stream per destination (several copy engines at once) --
      // measured on 8 GPUs 3.87 ms per step against 3.44 ms: the all-to-all-at-once pattern congests the receivers.
      multi_link = getenv("KH_CE_STREAMS") && atoi(getenv("KH_CE_STREAMS")) == 0;
      for (uint32_t q = 1; multi_link && q < p; ++q) {
        if (!x->link_stream[q]) KH_CUDA(cudaStreamCreateWithFlags(&x->link_stream[q], cudaStreamNonBlocking));
        if (!x->link_done[q]) KH_CUDA(cudaEventCreateWithFlags(&x->link_done[q], cudaEventDisableTiming));
      }
      if (!x->h_piece_counts) KH_CUDA(cudaHostAlloc((void **)&x->h_piece_counts, 16 * WIN_MAX_DST * sizeof(unsigned long long), cudaHostAllocDefault));
      C = x->copy_stream;
      stage.alloc((size_t)p * slab_cap * W);
      for (auto &e : evp) KH_CUDA(cudaEventCreateWithFlags(&e, ev_flags));
      cudaEvent_t e0;                                         // the copy stream starts behind what S has queued (the memsets above)
      KH_CUDA(cudaEventCreateWithFlags(&e0, cudaEventDisableTiming));
      KH_CUDA(cudaEventRecord(e0, S)); KH_CUDA(cudaStreamWaitEvent(C, e0, 0)); KH_CUDA(cudaEventDestroy(e0));
    }
    const uint64_t my_slab = ((uint64_t)parity * p + x->rank) * slab_cap;       // this rank's slab in every owner's buffer
    // the partition kernels leave one SM's worth of CTAs unlaunched, as the level-1 kernels of the pieces do (kh::bulk_feed): the
    // one-warp barrier kernels must find a free SM while those persistent kernels run
    const uint64_t ps_ctas = (uint64_t)std::max(t->sms - 1, 1) * 3;
    if (ce) {
      for (int c = 0; c < chunks; ++c) {                      // all partitions first: they depend on nothing
        const uint64_t k0 = n * (uint64_t)c / chunks, k1 = n * (uint64_t)(c + 1) / chunks, m = k1 - k0;
        unsigned long long *cursors = d_counts.p + (size_t)c * p;
        if (m) {
          BinDst dst{};
          for (uint32_t r = 0; r < p; ++r)
            dst.base[r] = r == x->rank ? x->local + (my_slab + (uint64_t)c * sub_cap) * W : stage.p + ((uint64_t)r * slab_cap + (uint64_t)c * sub_cap) * W;
          const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((m + tile_send - 1) / tile_send, ps_ctas));
          if (W == 1) KH_LAUNCH((scatter_slabs_kernel<1, false, false, false>), grid, PS_BLOCK, smem, S, d_keys + k0 * W, m, dist, t->hash_kind, t->hp, dst,
                                (unsigned long long)sub_cap, cursors, (unsigned long long *)nullptr, (uint32_t *)nullptr);
          else KH_LAUNCH((scatter_slabs_kernel<2, false, false, false>), grid, PS_BLOCK, smem, S, d_keys + k0 * W, m, dist, t->hash_kind, t->hp, dst,
                         (unsigned long long)sub_cap, cursors, (unsigned long long *)nullptr, (uint32_t *)nullptr);
        }
        KH_CUDA(cudaMemcpyAsync(x->h_piece_counts + (size_t)c * WIN_MAX_DST, cursors, p * sizeof(unsigned long long), cudaMemcpyDeviceToHost, S));
        KH_CUDA(cudaEventRecord(evp[c], S));
      }
    }
    for (int c = 0; c < chunks; ++c) {
      const uint64_t k0 = n * (uint64_t)c / chunks, k1 = n * (uint64_t)(c + 1) / chunks, m = k1 - k0;
      unsigned long long *cursors = d_counts.p + (size_t)c * p;
      if (ce) {
        KH_CUDA(cudaEventSynchronize(evp[c]));                // piece c is partitioned and its counts are here
        for (uint32_t q = 1; q < p; ++q) {                    // start with the right-hand neighbour: the ranks' copies do not pile up on one owner
          const uint32_t r = (x->rank + q) % p;
          const uint64_t cnt = std::min<uint64_t>(x->h_piece_counts[(size_t)c * WIN_MAX_DST + r], sub_cap);
          if (!cnt) continue;
          // link stream q is ordered behind its copy of the previous piece and behind nothing else: the host has seen piece c's
          // partition complete; the barrier stream waits for all of them
          KH_CUDA(cudaMemcpyAsync(x->remote[r] + (my_slab + (uint64_t)c * sub_cap) * W, stage.p + ((uint64_t)r * slab_cap + (uint64_t)c * sub_cap) * W,
                                  cnt * 8 * W, cudaMemcpyDeviceToDevice, multi_link ? x->link_stream[q] : C));
          if (multi_link) {
            KH_CUDA(cudaEventRecord(x->link_done[q], x->link_stream[q]));
            KH_CUDA(cudaStreamWaitEvent(C, x->link_done[q], 0));
          }
        }
      } else if (m) {
        BinDst dst{};
        for (uint32_t r = 0; r < p; ++r) dst.base[r] = x->remote[r] + (my_slab + (uint64_t)c * sub_cap) * W;
        // one CTA per SM: the level-1 kernel of the previous piece holds the rest of the SM
        const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((m + tile_send - 1) / tile_send, c == 0 ? ps_ctas : ps_ctas / 3));
        if (W == 1) KH_LAUNCH((scatter_slabs_kernel<1, false, false>), grid, PS_BLOCK, smem, S, d_keys + k0 * W, m, dist, t->hash_kind, t->hp, dst,
                              (unsigned long long)sub_cap, cursors, (unsigned long long *)nullptr, (uint32_t *)nullptr);
        else KH_LAUNCH((scatter_slabs_kernel<2, false, false>), grid, PS_BLOCK, smem, S, d_keys + k0 * W, m, dist, t->hash_kind, t->hp, dst,
                       (unsigned long long)sub_cap, cursors, (unsigned long long *)nullptr, (uint32_t *)nullptr);
      }
      const int slot = (int)(x->count_syncs & 1);
      ++x->count_syncs;
      KH_LAUNCH(peer_sync_kernel, 1, 32, 0, C, hd, x->rank, p, (unsigned long long)(epoch0 + c), cursors, 1, slot, d_flags.p, timeout_ns);
      KH_LAUNCH(make_segs_kernel, 1, 32, 0, C, hd.h[x->rank], slot, p, x->rank, (unsigned long long)sub_cap,
                (unsigned long long)((uint64_t)parity * p * slab_cap + (uint64_t)c * sub_cap), (unsigned long long)slab_cap,
                (unsigned long long)tile_l1, d_segs.p + c, d_matrix.p + (size_t)c * p * p, d_flags.p + 1);
      KH_CUDA(cudaEventRecord(ev[c], C));
      KH_CUDA(cudaStreamWaitEvent(R, ev[c], 0));
      bulk_pipe_feed(b, t, x->local, (uint64_t)p * sub_cap, d_segs.p + c, R);
      if (tr_on) KH_CUDA(cudaEventRecord(evt[1 + c], R));
    }
    std::vector<unsigned long long> h_matrix((size_t)chunks * p * p);
    int h_flags[2] = {0, 0};
    KH_CUDA(cudaMemcpyAsync(h_matrix.data(), d_matrix.p, h_matrix.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost, C));
    KH_CUDA(cudaMemcpyAsync(h_flags, d_flags.p, sizeof(h_flags), cudaMemcpyDeviceToHost, C));
    KH_CUDA(cudaStreamSynchronize(C));
    KH_CUDA(cudaStreamSynchronize(S));
    KH_CUDA(cudaStreamSynchronize(R));
    if (h_flags[0]) {
      x->poisoned = true;
      set_error("khx_insert_pipelined_dev: a peer did not reach the barrier in time; buffers are unusable");
      return KH_ERR_CUDA;
    }
    uint64_t n_recv = 0;
    for (uint32_t a = 0; a < p * p; ++a) counts_matrix[a] = 0;
    for (int c = 0; c < chunks; ++c)
      for (uint32_t a = 0; a < p; ++a)
        for (uint32_t r = 0; r < p; ++r) {
          const unsigned long long v = h_matrix[((size_t)c * p + a) * p + r];
          counts_matrix[a * p + r] += v;
          if (r == x->rank) n_recv += v;
        }
    *overflowed = h_flags[1] ? 1 : 0;
    if (h_flags[1] || n_recv == 0) return 0;                  // table untouched
    // the received keys as one array, should the general route have to take over
    b.refetch = [&](DevBuf<uint64_t> &flat, cudaStream_t s) -> const uint64_t * {
      flat.alloc(n_recv * W);
      uint64_t off = 0;
      for (int c = 0; c < chunks; ++c)
        for (uint32_t a = 0; a < p; ++a) {
          const uint64_t cnt = h_matrix[((size_t)c * p + a) * p + x->rank];
          if (cnt) cudaMemcpyAsync(flat.p + off * W, x->local + (((uint64_t)parity * p + a) * slab_cap + (uint64_t)c * sub_cap) * W, cnt * 8 * W,
                                   cudaMemcpyDeviceToDevice, s);
          off += cnt;
        }
      return flat.p;
    };
    const int64_t got = bulk_pipe_finish(b, t, n_recv, default_val, S);
    if (tr_on) {
      KH_CUDA(cudaEventRecord(evt[1 + chunks], S));
      KH_CUDA(cudaEventSynchronize(evt[1 + chunks]));
      auto at = [&](cudaEvent_t e) { float ms = 0; cudaEventElapsedTime(&ms, evt[0], e); return ms; };
      char line[512]; int o = snprintf(line, sizeof line, "[kh-trace] pipelined(%s, rank %u) ms since start:", ce ? "ce" : "sm", x->rank);
      for (int c = 0; c < chunks; ++c) {
        if (ce) o += snprintf(line + o, sizeof line - o, " P%d %.3f", c, at(evp[c]));
        o += snprintf(line + o, sizeof line - o, " B%d %.3f L%d %.3f |", c, at(ev[c]), c, at(evt[1 + c]));
      }
      snprintf(line + o, sizeof line - o, " done %.3f", at(evt[1 + chunks]));
      fprintf(stderr, "%s\n", line);
    }
    return got;
  });
}

int64_t khx_insert_pipelined_dev(kh_table *t, khx_peer *x, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n,
                                 uint64_t slab_cap, int parity, int chunks, uint64_t epoch0, uint32_t default_val,
                                 uint64_t *counts_matrix, int *overflowed) {
  return insert_pipelined_impl(t, x, dist_hash, dist_seed, d_keys, n, slab_cap, parity, chunks, epoch0, default_val, counts_matrix, overflowed, false);
}
int64_t khx_insert_ce_dev(kh_table *t, khx_peer *x, int dist_hash, uint32_t dist_seed, const uint64_t *d_keys, uint64_t n,
                          uint64_t slab_cap, int parity, int chunks, uint64_t epoch0, uint32_t default_val,
                          uint64_t *counts_matrix, int *overflowed) {
  return insert_pipelined_impl(t, x, dist_hash, dist_seed, d_keys, n, slab_cap, parity, chunks, epoch0, default_val, counts_matrix, overflowed, true);
}

}  // extern "C"
