// kh_window.cuh -- exact (histogram first) partition of a batch into <= 1024 bins: owner ranks of the multi-GPU exchange
// (partition.cu, exchange.cu; KH/distributed_batched_robinhood_map.hpp:632-741 assign_count_permute) or table regions.
//   1. bin_hist_kernel + bin_cursor_kernel: keys per bin -> start offsets ("cursors")
//   2. window_scatter_kernel: a block takes a tile of keys, counts them per bin in shared memory,
//      reserves its output ranges with ONE global atomicAdd per non-empty (tile, bin), sorts the
//      tile by bin in shared memory and writes each bin's run with coalesced stores.
// Traffic: 8W B read + 8W B written per key, streaming.  (The table's own bulk insert partitions with kh_bulk.cuh.)
#pragma once
#include <cstdint>
#include <cub/block/block_scan.cuh>
#include "kh_table.cuh"
#include "kh_hll.cuh"

namespace kh {

// what a key is binned by: mode 0 = table region (top `lp` bits of the spread STORE hash), mode 1 = owner rank
// of the multi-GPU exchange (DIST hash & (p-1) or % p, KH/distributed_batched_robinhood_map.hpp:164-178)
struct BinSpec {
  int mode, lp;
  uint32_t p;
  int hash_kind;
  HashParams hp;
};
template <int W>
__device__ __forceinline__ uint32_t bin_of(const BinSpec &b, const uint64_t (&k)[W]) {
  const uint64_t h = hash_runtime<W>(b.hash_kind, k, b.hp);
  if (b.mode == 0) return (uint32_t)top_bits(spread_hash(h), 64 - b.lp);
  return (b.p & (b.p - 1)) == 0 ? (uint32_t)(h & (uint64_t)(b.p - 1)) : (uint32_t)(h % (uint64_t)b.p);
}

constexpr int WIN_BLOCK = 256;
constexpr int WIN_TILE_BYTES = 16384;          // keys staged per tile
constexpr int WIN_MAX_P = REGION_FINE;         // 1024

// cursors[r] = number of keys in regions < r  (P = 2^lp regions, P <= 1024)
static __global__ void __launch_bounds__(REGION_FINE) window_cursor_kernel(const unsigned long long *fine, int lp,
                                                                    unsigned long long *cursors, int cstride = 1) {
  using Scan = cub::BlockScan<unsigned long long, REGION_FINE>;
  __shared__ typename Scan::TempStorage tmp;
  const int P = 1 << lp, per = REGION_FINE >> lp;
  unsigned long long c = 0;
  if ((int)threadIdx.x < P)
    for (int j = 0; j < per; ++j) c += fine[threadIdx.x * per + j];
  unsigned long long start;
  Scan(tmp).ExclusiveSum(c, start);
  if ((int)threadIdx.x < P) cursors[threadIdx.x * cstride] = start;
}

// keys per bin (n_bins <= 1024): shared-memory histogram per block, one global atomic per non-empty bin
template <int W>
__global__ void __launch_bounds__(512) bin_hist_kernel(const uint64_t *keys, uint64_t n, BinSpec spec, int n_bins,
                                                       unsigned long long *g_hist) {
  __shared__ uint32_t sh[WIN_MAX_P];
  for (int i = threadIdx.x; i < n_bins; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t k[W];
    load_key<W>(keys, i, k);
    atomicAdd(&sh[bin_of<W>(spec, k)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n_bins; i += blockDim.x)
    if (sh[i]) atomicAdd(&g_hist[i], (unsigned long long)sh[i]);
}

// cursors[b] = number of keys in bins < b; totals[b] (may be null) = keys in bin b
static __global__ void __launch_bounds__(REGION_FINE) bin_cursor_kernel(const unsigned long long *hist, int n_bins,
                                                                 unsigned long long *cursors, unsigned long long *totals) {
  using Scan = cub::BlockScan<unsigned long long, REGION_FINE>;
  __shared__ typename Scan::TempStorage tmp;
  const unsigned long long c = (int)threadIdx.x < n_bins ? hist[threadIdx.x] : 0ULL;
  unsigned long long start;
  Scan(tmp).ExclusiveSum(c, start);
  if ((int)threadIdx.x < n_bins) { cursors[threadIdx.x] = start; if (totals) totals[threadIdx.x] = c; }
}

// per-bin destination base pointers (possibly buffers of PEER GPUs mapped over NVLink): bin r's keys go to
// base[r][cursor ...], cursors counting from zero inside every bin
constexpr int WIN_MAX_DST = 8;
struct BinDst { uint64_t *base[WIN_MAX_DST]; };

// PERM: out_perm receives the original index of every output position (32-bit: n < 2^32)
// MULTI: write bin r through dst.base[r] instead of `out` (the fused partition + exchange of exchange.cu)
// TAGS (mode 0 only): out_tags receives, next to every sorted key, (spread hash >> tag_shift) & tag_mask = the key's
// home bucket index inside its region (kh_tile.cuh stages table tiles in shared memory and needs no hash of its own)
// cstride: distance (in elements) between two regions' cursors.  With ~1000 regions every tile reserves ~900 ranges;
// cursors packed 16 to a 128-byte line serialise those atomics in a few L2 slices, one line per cursor does not.
struct TagSpec { uint16_t *out; int shift; uint32_t mask; int cstride; };
template <int W, bool PERM, bool MULTI = false, bool TAGS = false>
__global__ void __launch_bounds__(WIN_BLOCK, 4) window_scatter_kernel(const uint64_t *keys, uint64_t n, BinSpec spec,
                                                                      unsigned long long *cursors, uint64_t *out,
                                                                      uint32_t *out_perm, BinDst dst = BinDst{},
                                                                      TagSpec tags = TagSpec{nullptr, 0, 0u, 1}) {
  constexpr int TILE = WIN_TILE_BYTES / (8 * W);      // keys per tile
  constexpr int KPT = TILE / WIN_BLOCK;               // keys per thread
  constexpr int RPT = WIN_MAX_P / WIN_BLOCK;          // regions per thread in the scan
  using Scan = cub::BlockScan<uint32_t, WIN_BLOCK>;
  __shared__ typename Scan::TempStorage tmp;
  __shared__ uint64_t skeys[TILE * W];
  __shared__ uint16_t sreg[TILE];
  __shared__ uint16_t ssrc[PERM ? TILE : 1];          // position of the key inside the tile before sorting
  __shared__ uint16_t stag[TAGS ? TILE : 1];
  __shared__ uint32_t cnt[WIN_MAX_P];                 // keys of the tile per region
  __shared__ uint32_t lo[WIN_MAX_P];                  // start of the region's run inside the sorted tile
  __shared__ unsigned long long gbase[WIN_MAX_P];     // start of the region's run in `out`
  const int P = spec.mode == 0 ? (1 << spec.lp) : (int)spec.p;
  const uint64_t n_tiles = (n + TILE - 1) / TILE;

  // keys of the NEXT tile are fetched into k[] as soon as the current tile's keys sit in shared memory, and the
  // range reservations (global atomics with return) are consumed only after the in-tile sort: both latencies
  // overlap shared-memory work instead of stalling the block between two barriers.
  uint64_t k[KPT][W];
  auto fetch = [&](uint64_t tile) {
    if (tile >= n_tiles) return;
    const uint64_t base = tile * TILE;
    const int tile_n = (int)min((uint64_t)TILE, n - base);
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
      const int e = j * WIN_BLOCK + threadIdx.x;
      if (e < tile_n) load_key<W>(keys, base + e, k[j]);
    }
  };
  fetch(blockIdx.x);

  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const uint64_t base = tile * TILE;
    const int tile_n = (int)min((uint64_t)TILE, n - base);
    for (int r = threadIdx.x; r < P; r += WIN_BLOCK) cnt[r] = 0;
    __syncthreads();

    uint32_t reg[KPT], loc[KPT], tg[TAGS ? KPT : 1];
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
      const int e = j * WIN_BLOCK + threadIdx.x;
      if (e < tile_n) {
        if constexpr (TAGS) {
          const uint64_t hm = spread_hash(hash_runtime<W>(spec.hash_kind, k[j], spec.hp));
          reg[j] = (uint32_t)top_bits(hm, 64 - spec.lp);
          tg[j] = (uint32_t)(hm >> tags.shift) & tags.mask;
        } else {
          reg[j] = bin_of<W>(spec, k[j]);
        }
        loc[j] = atomicAdd(&cnt[reg[j]], 1u);
      }
    }
    __syncthreads();

    // exclusive scan of cnt[0..P) -> lo ; reserve the output ranges
    uint32_t c[RPT], st[RPT];
    unsigned long long gb[RPT];
#pragma unroll
    for (int q = 0; q < RPT; ++q) { const int r = threadIdx.x * RPT + q; c[q] = r < P ? cnt[r] : 0u; }
    Scan(tmp).ExclusiveSum(c, st);
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int r = threadIdx.x * RPT + q;
      gb[q] = 0;
      if (r < P) {
        lo[r] = st[q];
        if (c[q]) gb[q] = atomicAdd(&cursors[TAGS ? r * tags.cstride : r], (unsigned long long)c[q]);
      }
    }
    __syncthreads();

    // sort the tile by region in shared memory
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
      const int e = j * WIN_BLOCK + threadIdx.x;
      if (e < tile_n) {
        const uint32_t pos = lo[reg[j]] + loc[j];
#pragma unroll
        for (int q = 0; q < W; ++q) skeys[pos * W + q] = k[j][q];
        sreg[pos] = (uint16_t)reg[j];
        if (PERM) ssrc[pos] = (uint16_t)e;
        if constexpr (TAGS) stag[pos] = (uint16_t)tg[j];
      }
    }
    fetch(tile + gridDim.x);
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int r = threadIdx.x * RPT + q;
      if (r < P && c[q]) gbase[r] = gb[q];
    }
    __syncthreads();

    // coalesced runs out
    for (int e = threadIdx.x; e < tile_n; e += WIN_BLOCK) {
      const uint32_t r = sreg[e];
      const unsigned long long gpos = gbase[r] + (uint32_t)(e - lo[r]);
      uint64_t *o = (MULTI ? dst.base[r] : out) + gpos * W;
#pragma unroll
      for (int q = 0; q < W; ++q) o[q] = skeys[e * W + q];
      if (PERM) out_perm[gpos] = (uint32_t)(base + ssrc[e]);
      if constexpr (TAGS) tags.out[gpos] = stag[e];
    }
    __syncthreads();
  }
}

}  // namespace kh
