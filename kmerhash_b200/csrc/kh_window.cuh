// kh_window.cuh -- region sort in front of the L2-windowed bulk insert (table.cu: insert_windowed).
//
// The batch is reordered so that keys of the same table region (top `lp` bits of the spread store hash
// = one contiguous 1/2^lp of the table, kh_table.cuh) are adjacent.  Order inside a region is free
// (COUNT is commutative), which allows the fast non-stable scheme:
//   1. fine histogram (1024 bins) -- produced by the HLL pass the insert runs anyway (kh_hll.cuh)
//   2. window_cursor_kernel: one block folds it into 2^lp region start offsets ("cursors")
//   3. window_scatter_kernel: a block takes a tile of keys, counts them per region in shared memory,
//      reserves its output ranges with ONE global atomicAdd per non-empty (tile, region), sorts the
//      tile by region in shared memory and writes each region's run with coalesced stores.
// Traffic: 8W B read + 8W B written per key, streaming.  This is the role the reference gives its
// radix-sort table flavour (KH/hashmap_radixsort.hpp: sort first so that the merge is sequential).
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

// ---- large-tile region scatter ------------------------------------------------------------------------------
// Same job as window_scatter_kernel<W,false,false,TAGS> (non-stable sort of the batch by table region, optional
// home-bucket tags), built for ~1000 regions: a CTA sorts RS_TILE_BYTES = 64 KB of keys at a time, so a region's
// run inside a tile is ~8 keys (64 B) instead of ~2 (16 B) -- 4x fewer, 4x longer global stores and 4x fewer range
// reservations (global atomics) per key.  Registers hold only the top 32 bits of every key's spread hash between
// the two sweeps (16 per thread); the key itself is read again from L2 in the second sweep instead of being kept:
//   sweep 1  load + hash, shared-memory histogram of the tile by region
//   scan     region starts inside the tile; ONE global atomicAdd per (tile, region) reserves the output range
//   sweep 2  re-read the key, take the next position of its region, store key / region / tag sorted in shared memory
//   out      sorted tile -> global memory, consecutive threads = consecutive addresses inside a run
constexpr int RS_BLOCK = 512;
constexpr int RS_TILE_BYTES = 65536;
// The batch may be handed over in up to WIN_MAX_DST pieces (the per-sender slabs of a peer receive buffer, exchange.cu):
// logical position i lies in piece s with prefix[s] <= i < prefix[s+1], at keys[base[s] + i - prefix[s]].
// tprefix[s] (filled by region_scatter_launch): kernel tiles in front of piece s -- every piece is cut into its own tiles,
// so a tile never straddles two pieces and no per-key address arithmetic is needed.
struct SegSpec {
  int nseg;
  unsigned long long prefix[WIN_MAX_DST + 1], base[WIN_MAX_DST], tprefix[WIN_MAX_DST + 1];
};
// sorted tile | per position: region (u16) + tag (u16), or one u32 (SLAB) | cur | goff (u32: positions stay < 2^32) |
// SLAB: 4096 HLL register bytes
template <int W> constexpr size_t rs_smem_bytes(bool slab = false) {
  return (size_t)RS_TILE_BYTES + 2 * (size_t)(RS_TILE_BYTES / (8 * W)) * 2 + WIN_MAX_P * 4 + WIN_MAX_P * 4 + (slab ? HLL_M : 0);
}
// SLAB mode: no histogram of the batch exists yet.  Region r owns the fixed range [r * slab, (r + 1) * slab) of `out` (the
// cursors start at r * slab), keys that do not fit go to an overflow list, and the SAME pass computes the HyperLogLog
// registers -- estimate, histogram and region sort in one read of the batch.  The tag is 32 bits wide: the top half of
// the spread hash (region = its top 10 bits), so the consumer can cut any table geometry out of it.
struct SlabSpec {
  unsigned long long slab;            // positions per region
  uint64_t *ovf_keys;                 // keys whose region was full
  unsigned long long *ovf_count, ovf_cap;
  uint32_t *g_regs;                   // HLL_M registers, merged with atomicMax
  uint32_t *tags32;                   // per output position
  int h32;                            // the store hash is 32 bits wide
};
template <int W, bool TAGS, bool SLAB = false>
__global__ void __launch_bounds__(RS_BLOCK, 2) region_scatter_kernel(const uint64_t *keys, uint64_t n, BinSpec spec,
                                                                     unsigned long long *cursors, uint64_t *out, TagSpec tags,
                                                                     const __grid_constant__ SegSpec segs, SlabSpec sl = SlabSpec{}) {
  constexpr int TILE = RS_TILE_BYTES / (8 * W);
  constexpr int KPT = TILE / RS_BLOCK;                // 16 (W=1) / 8 (W=2)
  constexpr int RPT = WIN_MAX_P / RS_BLOCK;           // 2
  constexpr int B = 8 / W;                            // keys loaded per thread before they are hashed
  using Scan = cub::BlockScan<uint32_t, RS_BLOCK>;
  __shared__ typename Scan::TempStorage tmp;
  extern __shared__ __align__(16) unsigned char rs_smem[];
  uint64_t *skeys = reinterpret_cast<uint64_t *>(rs_smem);
  uint32_t *goff = reinterpret_cast<uint32_t *>(rs_smem + RS_TILE_BYTES);      // run start in `out` - run start in tile (mod 2^32)
  uint32_t *cur = goff + WIN_MAX_P;                                            // histogram, then running position
  uint16_t *sreg = reinterpret_cast<uint16_t *>(cur + WIN_MAX_P);
  uint16_t *stag = sreg + TILE;
  uint32_t *shm = reinterpret_cast<uint32_t *>(sreg);                          // SLAB: hash top half per sorted position (replaces sreg + stag)
  uint32_t *regs = reinterpret_cast<uint32_t *>(sreg + 2 * TILE);              // SLAB: HLL register bytes, four to a word
  if constexpr (SLAB) {
    for (int i = threadIdx.x; i < HLL_M / 4; i += RS_BLOCK) regs[i] = 0;
  }
  const int P = 1 << spec.lp;
  const int rshift = spec.lp ? 32 - spec.lp : 31;     // lp == 0: one region (bit 31 is masked off below)
  const uint32_t rmask = spec.lp ? 0xffffffffu : 0u;
  const int tshift = tags.shift - 32;                 // tag = bits of the spread hash from tags.shift up
  const bool pieces = !SLAB && segs.nseg > 1;         // the one-read route is fed contiguous chunks
  const uint64_t n_tiles = pieces ? segs.tprefix[segs.nseg] : (n + TILE - 1) / TILE;
  // where tile t starts in `keys` and how many keys it holds (tile-uniform; a handful of compares per tile)
  auto locate = [&](uint64_t t, uint64_t &start, int &cnt) {
    if (!pieces) { start = t * TILE; cnt = (int)min((uint64_t)TILE, n - start); return; }
    int sg = 0;
#pragma unroll
    for (int q = 1; q < WIN_MAX_DST; ++q) sg += (q < segs.nseg && t >= segs.tprefix[q]) ? 1 : 0;
    const uint64_t off = (t - segs.tprefix[sg]) * TILE;
    start = segs.base[sg] + off;
    cnt = (int)min((uint64_t)TILE, (uint64_t)(segs.prefix[sg + 1] - segs.prefix[sg] - off));
  };

  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    uint64_t base;                                    // first key of the tile in `keys`
    int tile_n;
    locate(tile, base, tile_n);
    const uint64_t next = tile + gridDim.x;
    if (next < n_tiles) {                             // pull the next tile into L2 while this one is sorted
      uint64_t nb; int nn;
      locate(next, nb, nn);
      const int pe = (int)threadIdx.x * (128 / (8 * W));
      if (pe < nn) asm volatile("prefetch.global.L2 [%0];" ::"l"(keys + (nb + pe) * W));
    }
    for (int r = threadIdx.x; r < P; r += RS_BLOCK) cur[r] = 0;
    __syncthreads();

    uint32_t hm[KPT];
#pragma unroll
    for (int j0 = 0; j0 < KPT; j0 += B) {
      uint64_t k[B][W];
#pragma unroll
      for (int j = 0; j < B; ++j) {
        const int e = (j0 + j) * RS_BLOCK + threadIdx.x;
        if (e < tile_n) load_key<W>(keys, base + e, k[j]);
      }
#pragma unroll
      for (int j = 0; j < B; ++j) {
        const int e = (j0 + j) * RS_BLOCK + threadIdx.x;
        if (e < tile_n) {
          // the default store hash without the run-time dispatch
          const uint64_t h = spec.hash_kind == KH_HASH_MURMUR64AVX ? hash_key<KH_HASH_MURMUR64AVX, W>(k[j], spec.hp)
                                                                    : hash_runtime<W>(spec.hash_kind, k[j], spec.hp);
          hm[j0 + j] = (uint32_t)(spread_hash(h) >> 32);
          atomicAdd(&cur[(hm[j0 + j] >> rshift) & rmask], 1u);
          if constexpr (SLAB) {                       // hyperloglog64::update (KH/hyperloglog64.hpp:145-201), byte-wide max in a word
            uint32_t idx, rank;
            if (sl.h32) {
              const uint32_t x = (uint32_t)h;
              idx = x >> (32 - HLL_P);
              rank = __clz((int)((x << HLL_P) | 0xFFFu)) + 1;
            } else {
              idx = (uint32_t)(h >> (64 - HLL_P));
              rank = __clzll((long long)((h << HLL_P) | 0xFFFULL)) + 1;
            }
            uint32_t *w = regs + idx / 4;
            const int sh = (idx & 3) * 8;
            uint32_t old = *w;
            while (((old >> sh) & 0xffu) < rank) {
              const uint32_t seen = atomicCAS(w, old, (old & ~(0xffu << sh)) | (rank << sh));
              if (seen == old) break;
              old = seen;
            }
          }
        }
      }
    }
    __syncthreads();

    uint32_t c[RPT], st[RPT];
    unsigned long long gb[RPT];
#pragma unroll
    for (int q = 0; q < RPT; ++q) { const int r = threadIdx.x * RPT + q; c[q] = r < P ? cur[r] : 0u; }
    Scan(tmp).ExclusiveSum(c, st);
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int r = threadIdx.x * RPT + q;
      gb[q] = 0;
      if (r < P) {
        cur[r] = st[q];
        if (c[q]) gb[q] = atomicAdd(&cursors[(size_t)r * tags.cstride], (unsigned long long)c[q]);
      }
    }
    __syncthreads();

#pragma unroll
    for (int j0 = 0; j0 < KPT; j0 += B) {
      uint64_t k[B][W];
#pragma unroll
      for (int j = 0; j < B; ++j) {
        const int e = (j0 + j) * RS_BLOCK + threadIdx.x;
        if (e < tile_n) load_key<W>(keys, base + e, k[j]);
      }
#pragma unroll
      for (int j = 0; j < B; ++j) {
        const int e = (j0 + j) * RS_BLOCK + threadIdx.x;
        if (e < tile_n) {
          const uint32_t r = (hm[j0 + j] >> rshift) & rmask;
          const uint32_t pos = atomicAdd(&cur[r], 1u);
#pragma unroll
          for (int q = 0; q < W; ++q) skeys[pos * W + q] = k[j][q];
          if constexpr (SLAB) {
            shm[pos] = hm[j0 + j];
          } else {
            sreg[pos] = (uint16_t)r;
            if constexpr (TAGS) stag[pos] = (uint16_t)((hm[j0 + j] >> tshift) & tags.mask);
          }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < RPT; ++q) {
      const int r = threadIdx.x * RPT + q;
      if (r < P) goff[r] = (uint32_t)gb[q] - st[q];
    }
    __syncthreads();

#pragma unroll 4
    for (int e = threadIdx.x; e < tile_n; e += RS_BLOCK) {
      if constexpr (SLAB) {
        const uint32_t t32 = shm[e], r = (t32 >> rshift) & rmask;
        const uint32_t gpos = goff[r] + (uint32_t)e;
        if ((unsigned long long)gpos < (unsigned long long)(r + 1) * sl.slab) {
          uint64_t *o = out + (size_t)gpos * W;
#pragma unroll
          for (int q = 0; q < W; ++q) o[q] = skeys[e * W + q];
          sl.tags32[gpos] = t32;
        } else {                                      // the region's slab is full
          const unsigned long long at = atomicAdd(sl.ovf_count, 1ULL);
          if (at < sl.ovf_cap)
#pragma unroll
            for (int q = 0; q < W; ++q) sl.ovf_keys[at * W + q] = skeys[e * W + q];
        }
      } else {
        const size_t gpos = goff[sreg[e]] + (uint32_t)e;
        uint64_t *o = out + gpos * W;
#pragma unroll
        for (int q = 0; q < W; ++q) o[q] = skeys[e * W + q];
        if constexpr (TAGS) tags.out[gpos] = stag[e];
      }
    }
    __syncthreads();
  }
  if constexpr (SLAB) {
    for (int i = threadIdx.x; i < HLL_M; i += RS_BLOCK) {
      const uint32_t v = (regs[i / 4] >> ((i & 3) * 8)) & 0xffu;
      if (v) atomicMax(&sl.g_regs[i], v);
    }
  }
}

// SLAB mode: region r's cursor starts at r * slab
static __global__ void slab_cursor_kernel(unsigned long long *cursors, int cstride, unsigned long long slab) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < WIN_MAX_P) cursors[(size_t)r * cstride] = (unsigned long long)r * slab;
}

// host side: launch on `grid` CTAs (2 per SM fit)
template <int W, bool TAGS>
inline void region_scatter_launch(int sms, cudaStream_t s, const uint64_t *keys, uint64_t n, const BinSpec &spec,
                                  unsigned long long *cursors, uint64_t *out, const TagSpec &tags, const SegSpec *segs = nullptr) {
  // per device, so set on every launch (a process may drive several GPUs)
  KH_CUDA(cudaFuncSetAttribute(region_scatter_kernel<W, TAGS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs_smem_bytes<W>()));
  constexpr uint64_t TILE = RS_TILE_BYTES / (8 * W);
  SegSpec sg{};
  sg.nseg = 1;
  uint64_t n_tiles = (n + TILE - 1) / TILE;
  if (segs && segs->nseg > 1) {
    sg = *segs;
    sg.tprefix[0] = 0;
    for (int q = 0; q < sg.nseg; ++q) sg.tprefix[q + 1] = sg.tprefix[q] + (sg.prefix[q + 1] - sg.prefix[q] + TILE - 1) / TILE;
    n_tiles = sg.tprefix[sg.nseg];
  }
  const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>(n_tiles, (uint64_t)sms * 2));
  KH_LAUNCH((region_scatter_kernel<W, TAGS>), grid, RS_BLOCK, rs_smem_bytes<W>(), s, keys, n, spec, cursors, out, tags, sg);
}

// the cursors must have been set by slab_cursor_kernel; they persist over the launches of one batch (chunked feeding)
inline void region_scatter_slab_launch(int sms, cudaStream_t s, const uint64_t *keys, uint64_t n, const BinSpec &spec,
                                       unsigned long long *cursors, int cstride, uint64_t *out, const SlabSpec &sl) {
  KH_CUDA(cudaFuncSetAttribute(region_scatter_kernel<1, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rs_smem_bytes<1>(true)));
  constexpr uint64_t TILE = RS_TILE_BYTES / 8;
  const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + TILE - 1) / TILE, (uint64_t)sms * 2));
  SegSpec one{};
  one.nseg = 1;
  const TagSpec ts{nullptr, 32, 0u, cstride};
  KH_LAUNCH((region_scatter_kernel<1, true, true>), grid, RS_BLOCK, rs_smem_bytes<1>(true), s, keys, n, spec, cursors, out, ts, one, sl);
}

}  // namespace kh
