// kh_tile.cuh -- shared-memory-resident bulk insert ("tile insert") for COUNT tables with 8-byte keys.
//
// The L2-windowed insert (table.cu: insert_windowed) still pays one L2 atomic per key.  Here a contiguous
// piece of the table -- a TILE of TI_SLOTS slots = 128 KB of HBM -- is staged into the shared memory of one
// CTA (keys 64 KB + counts 32 KB; two CTAs per SM fit into B200's 227 KB), all keys whose home bucket lies
// in the tile are inserted with shared-memory atomics, and the tile is written back with coalesced 16-byte
// stores.  The table is read once and written once, streaming; no L2 atomic is issued at all.
//
// Input = the batch sorted by COARSE region (kh_window.cuh, 2^lp bins, lp <= 10) plus a 16-bit TAG per key
// (the key's home bucket index inside its coarse region, produced by the scatter pass that hashes the key
// anyway).  A coarse region spans SUBS = 2^subs_log2 tiles; the SUBS CTAs working on one region each scan the
// region's tags (2 B/key, L2 hits after the first CTA) and keep only the keys of their own tile: a warp
// compacts the matching positions into a small shared-memory queue and inserts them 32 at a time, so the
// probe code runs with full warps.
//
// Probe rule = the table's own (kh_table.cuh): 64-byte buckets of four slots, linear over buckets, a key is
// claimed in the FIRST empty slot in probe order (64-bit CAS in shared memory), a lost race re-scans.  A key
// whose chain would leave the tile (or the all-ones key, which lives in the side slot) is DEFERRED: its
// position goes to defer_idx and the host inserts those few keys through the global kernel afterwards.
// Preconditions checked by the host: W == 1, 64-byte buckets, COUNT reducer with a uniform addend, no
// sticky bits in the table (no erase since the last rebuild), aux words all zero.
#pragma once
#include <cstdint>
#include "kh_table.cuh"

namespace kh {

constexpr int TI_SLOTS = 8192;                  // slots per tile
constexpr int TI_BUCKETS = TI_SLOTS / 4;        // 64-byte buckets per tile
constexpr int TI_BUCKET_BITS = 11;
constexpr int TI_THREADS = 512;
constexpr int TI_WARPS = TI_THREADS / 32;
constexpr int TI_QUEUE = 160;                   // per-warp queue: 31 left over + 4 x 32 new
constexpr int TI_CHUNK = 128;                   // tags scanned per warp step (4 per lane, one 8-byte load)
constexpr size_t TI_SMEM = (size_t)TI_SLOTS * 8 + (size_t)TI_SLOTS * 4 + (size_t)TI_WARPS * TI_QUEUE * 4;

struct TileArgs {
  TableView t;
  const uint64_t *keys;                  // batch sorted by coarse region
  const uint16_t *tags;                  // home bucket inside the coarse region; padded to a multiple of TI_CHUNK
  const unsigned long long *bin_end;     // end offset of every coarse region in keys/tags (scatter cursors)
  int bin_stride;                        // elements between two regions' cursors
  int subs_log2;                         // tiles per coarse region = 1 << subs_log2
  int fresh;                             // the table holds no entry: skip reading it (2: slots uninitialised, write every tile)
  uint32_t dflt;                         // addend of every key
  unsigned long long *n_new;             // [0] += new keys, [1] += deferred keys
  uint32_t *defer_idx;                   // positions (in `keys`) of the deferred keys
  // SLAB layout (region_scatter_kernel<.., SLAB>): fine region f of 1024 owns positions [f * slab, bin_end[f]) and every
  // position carries the top 32 bits of its key's spread hash instead of a 16-bit tag
  const uint32_t *tags32;
  unsigned long long slab;
  int lbk;                               // log2(number of buckets of the table)
};

__device__ __forceinline__ void lds_v2(const uint64_t *p, uint64_t &a, uint64_t &b) {
  const uint32_t addr = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ld.shared.v2.u64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr) : "memory");
}

// one key into the staged tile; returns 1 when the key was new, defers when the chain leaves the tile.
// The slot index is computed with selects so that every lane of a batch reaches the same ATOMS instruction.
__device__ __forceinline__ uint32_t tile_insert_one(uint64_t *sk, uint32_t *sv, uint64_t key, uint32_t lb, uint32_t v,
                                                    uint32_t pos, const TileArgs &a) {
  if (key == EMPTY_WORD) lb = TI_BUCKETS;       // side-slot key: global path
  uint32_t is_new = 0;
  while (lb < (uint32_t)TI_BUCKETS) {
    uint64_t k0, k1, k2, k3;
    lds_v2(sk + lb * 4, k0, k1);
    lds_v2(sk + lb * 4 + 2, k2, k3);
    int m = k3 == key ? 3 : -1;
    m = k2 == key ? 2 : m; m = k1 == key ? 1 : m; m = k0 == key ? 0 : m;
    if (m < 0) {
      int e = k3 == EMPTY_WORD ? 3 : -1;
      e = k2 == EMPTY_WORD ? 2 : e; e = k1 == EMPTY_WORD ? 1 : e; e = k0 == EMPTY_WORD ? 0 : e;
      if (e < 0) { ++lb; continue; }
      const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long *>(&sk[lb * 4 + e]), (unsigned long long)EMPTY_WORD,
                                     (unsigned long long)key);
      if (old != EMPTY_WORD && old != key) continue;      // another key took the slot: scan the same bucket again
      is_new = old == EMPTY_WORD ? 1u : 0u;
      m = e;
    }
    atomicAdd(&sv[lb * 4 + m], v);
    return is_new;
  }
  a.defer_idx[atomicAdd(a.n_new + 1, 1ULL)] = pos;
  return 0;
}

template <bool SLAB>
__global__ void __launch_bounds__(TI_THREADS, 2) tile_insert_kernel(TileArgs a) {
  extern __shared__ __align__(16) unsigned char ti_smem[];
  uint64_t *sk = reinterpret_cast<uint64_t *>(ti_smem);
  uint32_t *sv = reinterpret_cast<uint32_t *>(sk + TI_SLOTS);
  uint32_t *sq = sv + TI_SLOTS + (threadIdx.x >> 5) * TI_QUEUE;     // this warp's queue
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint64_t n_tiles = a.t.capacity / TI_SLOTS;
  const uint32_t subs_mask = (1u << a.subs_log2) - 1u;
  uint32_t fresh = 0;

  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const uint64_t c = tile >> a.subs_log2;
    const uint32_t sub = (uint32_t)tile & subs_mask;
    // batch positions are 32-bit (the host keeps n < 2^32)
    uint32_t begin = 0, end = 0;
    // SLAB: the tile's keys lie in `pieces` consecutive fine regions starting at f0 (one region when the table has at
    // least 1024 tiles, else 1024 / n_tiles regions per tile); a key is the tile's when the top bits of its hash say so
    const int tile_bits = a.lbk - TI_BUCKET_BITS;                  // bits of a tile index
    const uint32_t pieces = SLAB ? (tile_bits >= 10 ? 1u : (1u << (10 - tile_bits))) : 1u;
    const uint32_t f0 = SLAB ? (tile_bits >= 10 ? (uint32_t)(tile >> (tile_bits - 10)) : (uint32_t)tile << (10 - tile_bits)) : 0u;
    if constexpr (!SLAB) {
      begin = c ? (uint32_t)a.bin_end[(c - 1) * a.bin_stride] : 0u;
      end = (uint32_t)a.bin_end[c * a.bin_stride];
      if (a.fresh == 1 && begin == end) continue;   // nothing to add to an empty tile (uniform per CTA)
    }
    uint4 *g = reinterpret_cast<uint4 *>(a.t.slots) + tile * TI_SLOTS;

    // ---- stage the tile -------------------------------------------------------------------
    if (a.fresh) {
      for (int i = threadIdx.x; i < TI_SLOTS; i += TI_THREADS) { sk[i] = EMPTY_WORD; sv[i] = 0u; }
    } else {
      for (int i = threadIdx.x; i < TI_SLOTS; i += TI_THREADS) {
        uint64_t w[2];
        LdCg<2>::ld(g + i, w);
        sk[i] = w[0]; sv[i] = (uint32_t)w[1];
      }
    }
    __syncthreads();

    // ---- scan the region's tags, compact the tile's keys, insert them with full warps -----
    // One batch of 32 keys is always "in flight": its key / tag loads were issued a step earlier, so the gather
    // latency overlaps the probes of the batch before it.
    uint32_t qn = 0;
    bool pending = false;
    uint64_t pkey = 0; uint32_t ptag = 0, ppos = 0;
    const int lb_shift = SLAB ? 32 - a.lbk : 0;                        // SLAB: bucket index = hash top half >> lb_shift
    for (uint32_t piece = 0; piece < pieces; ++piece) {
    if constexpr (SLAB) {
      const uint32_t f = f0 + piece;
      begin = (uint32_t)(f * a.slab);
      end = (uint32_t)min((unsigned long long)a.bin_end[(size_t)f * a.bin_stride], (unsigned long long)(f + 1) * a.slab);
    }
    const uint32_t first = begin & ~(uint32_t)(TI_CHUNK - 1);
    uint64_t base = (uint64_t)first + (uint64_t)warp * TI_CHUNK;       // 64-bit: may step past 2^32 at the end
    uint4 nxt = make_uint4(0u, 0u, 0u, 0u);
    auto load_tags = [&](uint64_t at) -> uint4 {
      if constexpr (SLAB) return __ldg(reinterpret_cast<const uint4 *>(a.tags32 + at) + lane);
      const uint2 v = __ldg(reinterpret_cast<const uint2 *>(a.tags + at) + lane);
      return make_uint4(v.x & 0xffffu, v.x >> 16, v.y & 0xffffu, v.y >> 16);
    };
    if (base < end) nxt = load_tags(base);
    for (; base < end; base += (uint64_t)TI_WARPS * TI_CHUNK) {
      const uint4 cur = nxt;
      const uint64_t nb = base + (uint64_t)TI_WARPS * TI_CHUNK;
      if (nb < end) nxt = load_tags(nb);
      const uint32_t tg[4] = {cur.x, cur.y, cur.z, cur.w};
      const uint32_t p0 = (uint32_t)base + (uint32_t)lane * 4;
      const bool interior = base >= begin && base + TI_CHUNK <= end;   // warp-uniform
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t p = p0 + j;
        bool mine = SLAB ? (uint32_t)((uint64_t)tg[j] >> (lb_shift + TI_BUCKET_BITS)) == (uint32_t)tile
                         : (tg[j] >> TI_BUCKET_BITS) == sub;
        if (!interior) mine = mine && p >= begin && p < end;
        const uint32_t bal = __ballot_sync(FULL_MASK, mine);
        if (mine) sq[qn + __popc(bal & lt_mask)] = p;
        qn += __popc(bal);
      }
      __syncwarp();
      uint32_t h = 0;
      while (qn - h >= 32u) {
        const uint32_t pos = sq[h + lane];
        const uint64_t key = __ldg(a.keys + pos);
        const uint32_t tag = SLAB ? (__ldg(a.tags32 + pos) >> lb_shift) : (uint32_t)__ldg(a.tags + pos);
        if (pending) fresh += tile_insert_one(sk, sv, pkey, ptag & (uint32_t)(TI_BUCKETS - 1), a.dflt, ppos, a);
        pkey = key; ptag = tag; ppos = pos; pending = true;
        h += 32;
      }
      if (h) {                                       // move the remainder (< 32 entries) to the front
        const uint32_t r = qn - h;
        const uint32_t v = lane < r ? sq[h + lane] : 0u;
        __syncwarp();
        if (lane < r) sq[lane] = v;
        qn = r;
      }
      __syncwarp();
    }
    }   // pieces
    if (pending) fresh += tile_insert_one(sk, sv, pkey, ptag & (uint32_t)(TI_BUCKETS - 1), a.dflt, ppos, a);
    if (lane < qn) {
      const uint32_t pos = sq[lane];
      const uint64_t key = __ldg(a.keys + pos);
      const uint32_t lb = (SLAB ? (__ldg(a.tags32 + pos) >> lb_shift) : (uint32_t)__ldg(a.tags + pos)) & (uint32_t)(TI_BUCKETS - 1);
      fresh += tile_insert_one(sk, sv, key, lb, a.dflt, pos, a);
    }
    __syncthreads();

    // ---- write the tile back ----------------------------------------------------------------
    for (int i = threadIdx.x; i < TI_SLOTS; i += TI_THREADS) {
      const uint64_t k = sk[i];
      g[i] = make_uint4((uint32_t)k, (uint32_t)(k >> 32), sv[i], 0u);
    }
    __syncthreads();
  }
  fresh = warp_sum(fresh);
  if (lane == 0 && fresh) atomicAdd(a.n_new, (unsigned long long)fresh);
}

}  // namespace kh
