// kh_bulk.cuh -- the streaming bulk insert: MSD partition of the batch down to TABLE TILES, then a shared-memory
// tile insert.  COUNT tables, 8- and 16-byte keys, any capacity.
//
// What it stands in for: insert_batch / insert_batch_by_hash of the reference's offsets table
// (KH/robinhood_offset_hashmap_ptr.hpp:1660-1869, 1876-2400), which hides two dependent cache misses per key behind
// software prefetch.  Random probes into a table larger than the 126 MB L2 run at the DRAM random-access rate on a GPU
// too (12-20 G read-modify-write/s measured, profiles/r01_probes_l2_resident.jsonl), far below what 6.5 TB/s of streaming
// bandwidth allows, so a large batch is turned into streaming work -- the role the reference gives its radix-sort table
// flavour (KH/hashmap_radixsort.hpp: sort first so that the merge is sequential):
//
//   level 1  l1_scatter_kernel   reads the batch ONCE: store hash -> HyperLogLog registers (the estimate insert() makes
//                                anyway, KH/robinhood_offset_hashmap_ptr.hpp:2512-2535) and a non-stable partition into
//                                2^lp REGIONS (lp <= 12) of the hash space = contiguous 1/2^lp of the table at ANY capacity
//                                (home bucket = top bits of the spread hash, kh_table.cuh).  No histogram first: region r
//                                owns the fixed range [r * slab, (r+1) * slab) of the output, the rare key that does not fit
//                                goes to an overflow list.  Output = 16W-byte RECORDS {key, tag, extra}; tag = top 32 bits of
//                                the spread hash, from which every later step cuts its bits -- nothing is hashed twice.
//   (estimate -> reserve: the capacity, hence the number of 128 KB tiles, is known only now)
//   level 2  l2_hist / l2_scatter only when the table has more tiles than level 1 made regions: every region is split
//                                into its tiles by the next tag bits (exact: tags histogram -> scan -> scatter).
//   tiles    tile_insert_kernel  one CTA stages one TILE (128 KB of table) in shared memory, streams the tile's records
//                                (one contiguous run: region == tile, or a few consecutive regions) with coalesced 16-byte
//                                loads, inserts with shared-memory CAS / atomicAdd, writes the tile back.  The table is
//                                read (not even that when it was empty) and written once; no L2 atomic is issued.
//   rebuild  tile_rebuild_kernel doubling / shrinking / sticky-bit cleaning as ONE streaming pass old table -> new table:
//                                slot order is region order, so new tile t is fed from a contiguous range of old slots.
//
// Data movement on the way in is a bulk asynchronous copy (cp.async.bulk -> UBLKCP, completion on an mbarrier): the next
// tile of keys lands in shared memory while the current one is sorted and written out.
#pragma once
#include <cstdint>
#include <cooperative_groups.h>
#include <cub/block/block_scan.cuh>
#include "kh_table.cuh"
#include "kh_hll.cuh"

namespace kh {

constexpr int BK_TILE_BYTES = 131072;          // bytes of table per tile
constexpr int BK_BUCKET_BITS = 11;             // 64-byte buckets per tile = 2048
constexpr int BK_BUCKETS = 1 << BK_BUCKET_BITS;
constexpr int BK_MAX_LP = 12;                  // level-2 tiles per region <= 4096; level-1 regions <= 4096 (8-byte keys)
constexpr int BK_MAX_BINS = 1 << BK_MAX_LP;
// level-1 regions: 4096 for 8-byte keys, 8192 for 16-byte keys -- their tiles hold half the slots, so a table of the same
// bytes has twice the tiles, and a run of ONE 32-byte record is still a full sector (16-byte records want two per run)
template <int W> struct BkL1 { static constexpr int MAX_LP = 11 + W; static constexpr int BINS = 1 << MAX_LP; };
inline int bk_l1_max_lp(int W) { return W == 1 ? BkL1<1>::MAX_LP : BkL1<2>::MAX_LP; }
constexpr int BK_CSTRIDE = 16;                 // one 128-byte line per region cursor (atomics of different regions never share a line)
template <int W> struct BkGeom {
  static constexpr int SLOTS = BK_TILE_BYTES / (16 * W);       // 8192 / 4096 slots per tile
  static constexpr int BUCKET_SLOTS = 4 / W;                   // 64-byte buckets
};

// The batch may be handed over in up to BK_MAX_SEG pieces (the per-sender slabs of a peer receive buffer, exchange.cu):
// logical position i lies in piece s with prefix[s] <= i < prefix[s+1], at keys[(base[s] + i - prefix[s]) * W].
// tprefix[s]: kernel tiles in front of piece s -- every piece is cut into its own tiles, so a tile never straddles two
// pieces and no per-key address arithmetic is needed.
constexpr int BK_MAX_SEG = 8;
struct BkSegs {
  int nseg;
  unsigned long long prefix[BK_MAX_SEG + 1], base[BK_MAX_SEG], tprefix[BK_MAX_SEG + 1];
};

// ---- mbarrier / bulk-copy helpers (sm_90+; SASS: SYNCS.*, UBLKCP) ---------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void *src, uint32_t bytes, uint32_t mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
  }
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// =====================================================================================================================
// level 1: hash + HyperLogLog + partition into 2^lp regions, one read of the batch
// =====================================================================================================================
struct L1Args {
  const uint64_t *keys;
  uint64_t n;
  int lp;                              // 1 <= lp <= BK_MAX_LP
  int hash_kind, h32;                  // store hash; h32: it is 32 bits wide (HLL rank in that width)
  HashParams hp;
  unsigned long long *cursors;         // [2^lp] x BK_CSTRIDE, start at r * slab, persist over the launches of one batch
  uint4 *recs;                         // [(slab << lp) * W]
  unsigned long long slab;             // positions per region
  uint64_t *ovf_keys;                  // keys whose region was full
  unsigned long long *ovf_count, ovf_cap;
  uint32_t *g_regs;                    // HLL_M registers of the table (cumulative over batches), merged with atomicMax
  unsigned long long pos_base;         // ORD: batch position of keys[0] (the records carry their key's position in the batch)
};

constexpr int L1_THREADS = 1024;
template <int W> struct L1Cfg {
  static constexpr int KPT = 8 / W;                            // keys per thread and tile
  static constexpr int TILE = L1_THREADS * KPT;                // 8192 / 4096 keys = 64 KB
  static constexpr int RAW_BYTES = TILE * 8 * W;
  // raw tile | sorted keys | sorted tags | packed 16-bit counters | run offsets | HLL register bytes | mbarrier
  static constexpr int BINS = BkL1<W>::BINS;
  static constexpr int BPT = BINS / L1_THREADS;                // bins per thread in the run-start step: 4 / 8
  static constexpr size_t SMEM = (size_t)2 * RAW_BYTES + (size_t)TILE * 4 + BINS * 2 + BINS * 4 + HLL_M + 16;
  static constexpr size_t SMEM_ORD = SMEM + (size_t)TILE * 2;  // + where in the tile every sorted key came from
};

// dsegs: the pieces as a DEVICE-resident table (may be null: the launch argument `segs_arg` holds them).  The pipelined
// multi-GPU receiver (exchange.cu) launches this kernel before the host knows how many keys every sender stored: a small
// kernel behind the flag barrier writes the table, the grid is fixed and every CTA reads its tile count from it.
// ORD: the order-dependent reducers (FIRST / LAST with per-key values): every record also carries the position of its key in
// the batch (the spare word of the record), from which the tile insert decides the winner and fetches its value.
template <int W, bool ORD = false>
__global__ void __launch_bounds__(L1_THREADS, 1) l1_scatter_kernel(const L1Args a, const __grid_constant__ BkSegs segs_arg,
                                                                   const BkSegs *dsegs) {
  using Cfg = L1Cfg<W>;
  __shared__ BkSegs segs;
  {
    static_assert(sizeof(BkSegs) % 8 == 0, "BkSegs is copied word by word");
    const unsigned long long *src = reinterpret_cast<const unsigned long long *>(dsegs ? dsegs : &segs_arg);
    unsigned long long *dst = reinterpret_cast<unsigned long long *>(&segs);
    for (int i = threadIdx.x; i < (int)(sizeof(BkSegs) / 8); i += L1_THREADS) dst[i] = src[i];
    __syncthreads();
  }
  constexpr int TILE = Cfg::TILE, KPT = Cfg::KPT;
  using Scan = cub::BlockScan<uint32_t, L1_THREADS>;
  __shared__ typename Scan::TempStorage scan_tmp;
  extern __shared__ __align__(128) unsigned char l1_smem[];
  uint64_t *raw = reinterpret_cast<uint64_t *>(l1_smem);
  uint64_t *skeys = reinterpret_cast<uint64_t *>(l1_smem + Cfg::RAW_BYTES);
  uint32_t *stags = reinterpret_cast<uint32_t *>(l1_smem + 2 * Cfg::RAW_BYTES);
  uint32_t *cnt32 = stags + TILE;                              // two 16-bit counters per word: count, then run start
  constexpr int BINS = Cfg::BINS, BPT = Cfg::BPT;
  uint32_t *goff = cnt32 + BINS / 2;                           // run start in recs - run start in tile (mod 2^32)
  uint32_t *regs = goff + BINS;                                // HLL register bytes, four to a word
  const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(regs + HLL_M / 4);
  uint16_t *sorg = reinterpret_cast<uint16_t *>(regs + HLL_M / 4 + 4);   // ORD: [TILE] index inside the tile of every sorted key
  const uint32_t raw_s = (uint32_t)__cvta_generic_to_shared(raw);
  const int tid = threadIdx.x;

  for (int i = tid; i < HLL_M / 4; i += L1_THREADS) regs[i] = 0;
#pragma unroll
  for (int q = 0; q < BPT / 2; ++q) cnt32[(BPT / 2) * tid + q] = 0;
  if (tid == 0) { mbar_init(mbar, 1); fence_proxy_async(); }

  const int rshift = 32 - a.lp;
  const bool pieces = segs.nseg > 1 || dsegs != nullptr;
  const uint64_t n_tiles = pieces ? segs.tprefix[segs.nseg] : (a.n + TILE - 1) / TILE;
  // where tile t starts in `keys` and how many keys it holds (tile-uniform; a handful of compares per tile)
  auto locate = [&](uint64_t t, uint64_t &start, int &cnt) {
    if (!pieces) { start = segs.base[0] + t * TILE; cnt = (int)min((uint64_t)TILE, a.n - t * TILE); return; }
    int sg = 0;
#pragma unroll
    for (int q = 1; q < BK_MAX_SEG; ++q) sg += (q < segs.nseg && t >= segs.tprefix[q]) ? 1 : 0;
    const uint64_t off = (t - segs.tprefix[sg]) * TILE;
    start = segs.base[sg] + off;
    cnt = (int)min((uint64_t)TILE, (uint64_t)(segs.prefix[sg + 1] - segs.prefix[sg] - off));
  };
  // keys of tile t -> raw: one bulk asynchronous copy when source and size are 16-byte aligned (returns true: wait on the
  // mbarrier), else a cooperative copy that the next __syncthreads() publishes
  auto stage = [&](uint64_t t) -> bool {
    uint64_t start; int cnt;
    locate(t, start, cnt);
    const uint64_t *src = a.keys + start * W;
    const uint32_t bytes = (uint32_t)cnt * 8u * W;
    const bool bulk = (reinterpret_cast<uintptr_t>(src) & 15u) == 0 && bytes >= 16u;
    if (bulk) {
      const uint32_t body = bytes & ~15u;
      if (tid == 0) {
        fence_proxy_async();
        mbar_expect_tx(mbar, body);
        bulk_g2s(raw_s, src, body, mbar);
        if (body != bytes) raw[cnt - 1] = __ldg(src + cnt - 1);          // W == 1, odd count: the last key by hand
      }
    } else {
      for (int e = tid; e < cnt * W; e += L1_THREADS) raw[e] = __ldg(src + e);
    }
    return bulk;
  };

  __syncthreads();                                   // mbarrier initialised, counters zero
  uint32_t parity = 0;
  bool cur_bulk = false;
  if (blockIdx.x < n_tiles) cur_bulk = stage(blockIdx.x);
  __syncthreads();

  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    uint64_t base; int tile_n;
    locate(tile, base, tile_n);
    if (cur_bulk) { mbar_wait(mbar, parity); parity ^= 1u; }

    // ---- sweep 1: hash, HLL registers, rank inside the region's run of this tile -----------------------------------
    uint32_t hm[KPT], rk[KPT];
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
      const int e = j * L1_THREADS + tid;
      hm[j] = 0; rk[j] = 0;
      if (e < tile_n) {
        uint64_t k[W];
#pragma unroll
        for (int q = 0; q < W; ++q) k[q] = raw[e * W + q];
        const uint64_t h = hash_runtime<W>(a.hash_kind, k, a.hp);
        hm[j] = (uint32_t)(spread_hash(h) >> 32);
        const uint32_t r = hm[j] >> rshift;
        const int sh = (r & 1) << 4;
        rk[j] = (atomicAdd(&cnt32[r >> 1], 1u << sh) >> sh) & 0xffffu;
        // hyperloglog64::update (KH/hyperloglog64.hpp:145-201), byte-wide max inside a shared-memory word
        uint32_t idx, rank;
        if (a.h32) {
          const uint32_t x = (uint32_t)h;
          idx = x >> (32 - HLL_P);
          rank = __clz((int)((x << HLL_P) | 0xFFFu)) + 1;
        } else {
          idx = (uint32_t)(h >> (64 - HLL_P));
          rank = __clzll((long long)((h << HLL_P) | 0xFFFULL)) + 1;
        }
        uint32_t *w = regs + (idx >> 2);
        const int bs = (idx & 3) * 8;
        uint32_t old = *w;
        while (((old >> bs) & 0xffu) < rank) {
          const uint32_t seen = atomicCAS(w, old, (old & ~(0xffu << bs)) | (rank << bs));
          if (seen == old) break;
          old = seen;
        }
      }
    }
    __syncthreads();

    // ---- run starts inside the tile; ONE global atomicAdd per (tile, region) reserves the output run ---------------
    uint32_t c[BPT], lo[BPT];
    uint32_t gb[BPT];                                  // record positions fit 32 bits (the host bounds the batch)
    {
      uint32_t w[BPT / 2];
#pragma unroll
      for (int q = 0; q < BPT / 2; ++q) w[q] = cnt32[(BPT / 2) * tid + q];
      uint32_t tot = 0, st;
#pragma unroll
      for (int q = 0; q < BPT; ++q) { c[q] = (w[q >> 1] >> ((q & 1) << 4)) & 0xffffu; tot += c[q]; }
      Scan(scan_tmp).ExclusiveSum(tot, st);
#pragma unroll
      for (int q = 0; q < BPT; ++q) { lo[q] = st; st += c[q]; }
#pragma unroll
      for (int q = 0; q < BPT / 2; ++q) cnt32[(BPT / 2) * tid + q] = lo[2 * q] | (lo[2 * q + 1] << 16);
#pragma unroll
      for (int q = 0; q < BPT; ++q) {
        gb[q] = 0;
        if (c[q]) gb[q] = (uint32_t)atomicAdd(&a.cursors[(size_t)(BPT * tid + q) * BK_CSTRIDE], (unsigned long long)c[q]);
      }
    }
    __syncthreads();

    // ---- sweep 2: keys and tags to their sorted position ------------------------------------------------------------
#pragma unroll
    for (int j = 0; j < KPT; ++j) {
      const int e = j * L1_THREADS + tid;
      if (e < tile_n) {
        const uint32_t r = hm[j] >> rshift;
        const uint32_t pos = ((cnt32[r >> 1] >> ((r & 1) << 4)) & 0xffffu) + rk[j];
#pragma unroll
        for (int q = 0; q < W; ++q) skeys[pos * W + q] = raw[e * W + q];
        stags[pos] = hm[j];
        if (ORD) sorg[pos] = (uint16_t)e;
      }
    }
#pragma unroll
    for (int q = 0; q < BPT; ++q) goff[BPT * tid + q] = gb[q] - lo[q];
    __syncthreads();

    // ---- the raw buffer is free: fetch the next tile while this one goes out ----------------------------------------
    const uint64_t next = tile + gridDim.x;
    if (next < n_tiles) cur_bulk = stage(next);
#pragma unroll
    for (int q = 0; q < BPT / 2; ++q) cnt32[(BPT / 2) * tid + q] = 0;

    // ---- sorted tile out: consecutive threads = consecutive records inside a run ------------------------------------
#pragma unroll 2
    for (int e = tid; e < tile_n; e += L1_THREADS) {
      const uint32_t t32 = stags[e], r = t32 >> rshift;
      const uint32_t gpos = goff[r] + (uint32_t)e;
      if ((unsigned long long)gpos < (unsigned long long)(r + 1) * a.slab) {
        uint4 *o = a.recs + (size_t)gpos * W;
        if (W == 1) {
          const uint64_t k = skeys[e];
          o[0] = make_uint4((uint32_t)k, (uint32_t)(k >> 32), t32, ORD ? (uint32_t)(a.pos_base + base + sorg[e]) : 0u);
        } else {
          const ulonglong2 k = *reinterpret_cast<const ulonglong2 *>(skeys + 2 * e);
          o[0] = make_uint4((uint32_t)k.x, (uint32_t)(k.x >> 32), (uint32_t)k.y, (uint32_t)(k.y >> 32));
          o[W - 1] = make_uint4(t32, ORD ? (uint32_t)(a.pos_base + base + sorg[e]) : 0u, 0u, 0u);
        }
      } else {                                        // the region's slab is full
        const unsigned long long at = atomicAdd(a.ovf_count, 1ULL);
        if (at < a.ovf_cap)
#pragma unroll
          for (int q = 0; q < W; ++q) a.ovf_keys[at * W + q] = skeys[e * W + q];
      }
    }
    __syncthreads();
  }

  for (int i = tid; i < HLL_M; i += L1_THREADS) {
    const uint32_t v = (regs[i >> 2] >> ((i & 3) * 8)) & 0xffu;
    if (v) atomicMax(&a.g_regs[i], v);
  }
}

// keys per level-1 tile (the pieces of a segmented batch are cut into tiles of this size)
inline uint64_t l1_tile_keys(int W) { return W == 1 ? L1Cfg<1>::TILE : L1Cfg<2>::TILE; }

// region r's cursor starts at r * slab
static __global__ void l1_cursor_kernel(unsigned long long *cursors, int n_regions, unsigned long long slab) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n_regions) cursors[(size_t)r * BK_CSTRIDE] = (unsigned long long)r * slab;
}

// =====================================================================================================================
// level 2: split every region into its T = 2^tbits tiles by the next tag bits (exact)
// =====================================================================================================================
constexpr int L2_THREADS = 512;
constexpr int L2_KPT = 8;
constexpr int L2_TILE = L2_THREADS * L2_KPT;       // records per chunk

struct L2Args {
  const uint4 *recs;                   // level-1 output
  const unsigned long long *cursors;   // level-1 cursors (end of every region's run, may exceed the slab: clamp)
  unsigned long long slab;
  int lp, tbits;                       // regions = 2^lp, tiles per region = 2^tbits (tbits <= 12)
  unsigned long long *chunk_prefix;    // [2^lp + 1] chunks of L2_TILE records in front of region r
  unsigned long long *tile_cnt;        // [2^(lp+tbits)] histogram, then (after the scan) unused (exact mode only)
  unsigned long long *tile_cur;        // [2^(lp+tbits)] x cstride: starts at tile_begin, ends as tile_end
  uint4 *recs2;                        // output: exact-packed, or one slab of slab2 positions per tile
  // slab mode (slab2 != 0): no histogram pass -- tile x owns [x * slab2, (x+1) * slab2), sized for a uniform hash like the
  // level-1 slabs; the rare record that does not fit gives its key to the overflow list.  cstride: words between two cursors.
  unsigned long long slab2;
  int cstride;
  uint64_t *ovf_keys;
  unsigned long long *ovf_count, ovf_cap;
};

__device__ __forceinline__ unsigned long long l2_region_len(const L2Args &a, uint32_t r) {
  const unsigned long long b = (unsigned long long)r * a.slab;
  unsigned long long e = a.cursors[(size_t)r * BK_CSTRIDE];
  e = min(e, b + a.slab);
  return e - b;
}

// chunk_prefix: one block, 1024 threads, 4 regions per thread
static __global__ void __launch_bounds__(1024) l2_plan_kernel(L2Args a) {
  using Scan = cub::BlockScan<unsigned long long, 1024>;
  __shared__ typename Scan::TempStorage tmp;
  const uint32_t R = 1u << a.lp;
  unsigned long long c[4], s[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t r = threadIdx.x * 4 + q;
    c[q] = r < R ? (l2_region_len(a, r) + L2_TILE - 1) / L2_TILE : 0ULL;
  }
  Scan(tmp).ExclusiveSum(c, s);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t r = threadIdx.x * 4 + q;
    if (r < R) a.chunk_prefix[r] = s[q];
    if (r == R - 1) a.chunk_prefix[R] = s[q] + c[q];
  }
}

// chunk c -> (region, first record, record count); binary search over chunk_prefix (block-uniform)
__device__ __forceinline__ void l2_locate(const L2Args &a, unsigned long long c, uint32_t &r, unsigned long long &first, int &cnt) {
  uint32_t lo = 0, hi = 1u << a.lp;                  // invariant: prefix[lo] <= c < prefix[hi]
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (a.chunk_prefix[mid] <= c) lo = mid; else hi = mid;
  }
  r = lo;
  const unsigned long long off = (c - a.chunk_prefix[r]) * L2_TILE;
  first = (unsigned long long)r * a.slab + off;
  cnt = (int)min((unsigned long long)L2_TILE, l2_region_len(a, r) - off);
}

template <int W>
__device__ __forceinline__ uint32_t rec_tag(const uint4 *recs, unsigned long long i) {
  const uint32_t *p = reinterpret_cast<const uint32_t *>(recs + i * W);
  return __ldg(p + (W == 1 ? 2 : 4));
}

template <int W>
__global__ void __launch_bounds__(L2_THREADS) l2_hist_kernel(L2Args a) {
  __shared__ uint32_t hist[BK_MAX_BINS];
  const uint32_t T = 1u << a.tbits;
  const int tshift = 32 - a.lp - a.tbits;
  const unsigned long long n_chunks = a.chunk_prefix[1u << a.lp];
  for (unsigned long long ch = blockIdx.x; ch < n_chunks; ch += gridDim.x) {
    uint32_t r; unsigned long long first; int cnt;
    l2_locate(a, ch, r, first, cnt);
    for (uint32_t i = threadIdx.x; i < T; i += L2_THREADS) hist[i] = 0;
    __syncthreads();
    for (int e = threadIdx.x; e < cnt; e += L2_THREADS)
      atomicAdd(&hist[(rec_tag<W>(a.recs, first + e) >> tshift) & (T - 1)], 1u);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < T; i += L2_THREADS)
      if (hist[i]) atomicAdd(&a.tile_cnt[((size_t)r << a.tbits) + i], (unsigned long long)hist[i]);
    __syncthreads();
  }
}

template <int W>
__global__ void __launch_bounds__(L2_THREADS, 2) l2_scatter_kernel(L2Args a) {
  using Scan = cub::BlockScan<uint32_t, L2_THREADS>;
  __shared__ typename Scan::TempStorage scan_tmp;
  extern __shared__ __align__(16) unsigned char l2_smem[];
  uint4 *srec = reinterpret_cast<uint4 *>(l2_smem);                        // [L2_TILE * W] sorted records
  uint32_t *cnt32 = reinterpret_cast<uint32_t *>(srec + L2_TILE * W);      // packed 16-bit counters
  uint32_t *goff = cnt32 + BK_MAX_BINS / 2;
  const int tid = threadIdx.x;
  const uint32_t T = 1u << a.tbits;
  const int tshift = 32 - a.lp - a.tbits;
  constexpr int BPT = BK_MAX_BINS / L2_THREADS;                            // 8 bins = 4 words per thread
  const unsigned long long n_chunks = a.chunk_prefix[1u << a.lp];
  for (unsigned long long ch = blockIdx.x; ch < n_chunks; ch += gridDim.x) {
    uint32_t r; unsigned long long first; int cnt;
    l2_locate(a, ch, r, first, cnt);
#pragma unroll
    for (int q = 0; q < BPT / 2; ++q) cnt32[(BPT / 2) * tid + q] = 0;
    __syncthreads();
    uint32_t sub[L2_KPT], rk[L2_KPT];
#pragma unroll
    for (int j = 0; j < L2_KPT; ++j) {
      const int e = j * L2_THREADS + tid;
      sub[j] = 0; rk[j] = 0;
      if (e < cnt) {
        sub[j] = (rec_tag<W>(a.recs, first + e) >> tshift) & (T - 1);
        const int sh = (sub[j] & 1) << 4;
        rk[j] = (atomicAdd(&cnt32[sub[j] >> 1], 1u << sh) >> sh) & 0xffffu;
      }
    }
    __syncthreads();
    uint32_t c[BPT], lo[BPT];
    unsigned long long gb[BPT];
    {
      const uint4 w4 = *reinterpret_cast<const uint4 *>(cnt32 + 4 * tid);
      c[0] = w4.x & 0xffffu; c[1] = w4.x >> 16; c[2] = w4.y & 0xffffu; c[3] = w4.y >> 16;
      c[4] = w4.z & 0xffffu; c[5] = w4.z >> 16; c[6] = w4.w & 0xffffu; c[7] = w4.w >> 16;
      uint32_t tot = 0, st;
#pragma unroll
      for (int q = 0; q < BPT; ++q) tot += c[q];
      Scan(scan_tmp).ExclusiveSum(tot, st);
#pragma unroll
      for (int q = 0; q < BPT; ++q) { lo[q] = st; st += c[q]; }
      *reinterpret_cast<uint4 *>(cnt32 + 4 * tid) = make_uint4(lo[0] | (lo[1] << 16), lo[2] | (lo[3] << 16), lo[4] | (lo[5] << 16), lo[6] | (lo[7] << 16));
#pragma unroll
      for (int q = 0; q < BPT; ++q) {
        gb[q] = 0;
        if (c[q]) gb[q] = atomicAdd(&a.tile_cur[(((size_t)r << a.tbits) + BPT * tid + q) * a.cstride], (unsigned long long)c[q]);
      }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < L2_KPT; ++j) {
      const int e = j * L2_THREADS + tid;
      if (e < cnt) {
        const uint32_t pos = ((cnt32[sub[j] >> 1] >> ((sub[j] & 1) << 4)) & 0xffffu) + rk[j];
#pragma unroll
        for (int q = 0; q < W; ++q) srec[pos * W + q] = __ldg(a.recs + (first + e) * W + q);
      }
    }
#pragma unroll
    for (int q = 0; q < BPT; ++q) goff[BPT * tid + q] = gb[q] - lo[q];
    __syncthreads();
    for (int e = tid; e < cnt; e += L2_THREADS) {
      const uint4 r0 = srec[e * W];
      const uint32_t t32 = W == 1 ? r0.z : srec[e * W + W - 1].x;
      const uint32_t sub = (t32 >> tshift) & (T - 1);
      const uint32_t gpos = goff[sub] + (uint32_t)e;
      if (a.slab2 == 0 || (unsigned long long)gpos < ((((unsigned long long)r << a.tbits) + sub + 1ULL) * a.slab2)) {
        a.recs2[(size_t)gpos * W] = r0;
        if (W == 2) a.recs2[(size_t)gpos * W + W - 1] = srec[e * W + W - 1];
      } else {                                          // the tile's slab is full
        const unsigned long long at = atomicAdd(a.ovf_count, 1ULL);
        if (at < a.ovf_cap) {
          a.ovf_keys[at * W] = (uint64_t)r0.x | ((uint64_t)r0.y << 32);
          if (W == 2) a.ovf_keys[at * W + W - 1] = (uint64_t)r0.z | ((uint64_t)r0.w << 32);
        }
      }
    }
    __syncthreads();
  }
}
constexpr size_t l2_smem_bytes(int W) { return (size_t)L2_TILE * 16 * W + BK_MAX_BINS * 2 + BK_MAX_BINS * 4; }

// =====================================================================================================================
// tile insert
// =====================================================================================================================
constexpr int TI_THREADS = 512;
template <int W> struct TiMlp { static constexpr int value = 4 / W; };   // records in flight per thread

struct TileArgs {
  TableView t;
  const uint4 *recs;
  // SLAB layout (straight from level 1): tile `x` owns the rpt = 2^rpt_log2 consecutive regions x * rpt ..., region r's
  // records are [r * slab, min(ends[r * BK_CSTRIDE], (r+1) * slab)).  EXACT layout (after level 2, slab == 0): tile x's
  // records are [x ? ends[x-1] : 0, ends[x]).
  const unsigned long long *ends;
  unsigned long long slab;
  int rpt_log2;
  int lbk;                               // log2(number of buckets of the table)
  int fresh;                             // 0: read the tile; 1: the table is empty and initialised (tiles without keys are
                                         // skipped); 2: slots uninitialised, every tile is written
  uint32_t dflt;                         // addend of every key
  unsigned long long *n_new;             // [0] += new keys, [1] += deferred keys
  uint32_t *defer_idx;                   // positions (in recs) of the deferred keys
  const uint32_t *vals;                  // ORD kernels: the batch's values by batch position (null: dflt for every key)
};

__device__ __forceinline__ void lds_v2(const uint64_t *p, uint64_t &a, uint64_t &b) {
  const uint32_t addr = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("ld.shared.v2.u64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr) : "memory");
}
// 128-bit compare-and-swap in shared memory (SASS: ATOMS.CAS.128)
__device__ __forceinline__ void cas128_shared(uint64_t *p, uint64_t e0, uint64_t e1, uint64_t d0, uint64_t d1, uint64_t &o0, uint64_t &o1) {
  const uint32_t addr = (uint32_t)__cvta_generic_to_shared(p);
  asm volatile("{\n\t.reg .b128 c, n, o;\n\tmov.b128 c, {%3, %4};\n\tmov.b128 n, {%5, %6};\n\t"
               "atom.relaxed.cta.shared::cta.cas.b128 o, [%2], c, n;\n\tmov.b128 {%0, %1}, o;\n\t}"
               : "=l"(o0), "=l"(o1) : "r"(addr), "l"(e0), "l"(e1), "l"(d0), "l"(d1) : "memory");
}

// One key into the staged tile.  Probe rule = the table's own (kh_table.cuh): 64-byte buckets, linear over buckets, a key is
// claimed in the FIRST empty slot in probe order, a lost race re-scans.  Returns 1 when the key was new, 0 when it was
// present, 2 when its chain leaves the tile (or it is the all-ones key of the side slot): the caller defers it.
// MODE: what happens to the value word of the key's slot: TP_ADD (COUNT), TP_SET (rebuild: keys are unique), TP_MIN / TP_MAX
// (FIRST / LAST: v is the encoded batch position, the smallest / largest wins; under TP_MIN a key that was in the table
// before the batch -- its bit in `oldbits` -- keeps its value word untouched).
enum : int { TP_ADD = 0, TP_SET = 1, TP_MIN = 2, TP_MAX = 3 };
template <int MODE>
__device__ __forceinline__ void tile_value(uint32_t *sv, uint32_t slot, uint32_t v, uint32_t is_new, const uint32_t *oldbits) {
  if (MODE == TP_SET) sv[slot] = v;
  else if (MODE == TP_ADD) atomicAdd(&sv[slot], v);
  else if (MODE == TP_MAX) atomicMax(&sv[slot], v);
  else if (is_new || !oldbits || !((oldbits[slot >> 5] >> (slot & 31)) & 1u)) atomicMin(&sv[slot], v);
}
template <int W, int MODE>
__device__ __forceinline__ uint32_t tile_put(uint64_t *sk, uint32_t *sv, const uint64_t (&key)[W], uint32_t lb, uint32_t v,
                                             const uint32_t *oldbits = nullptr) {
  if (key_is_empty<W>(key)) return 2u;
  while (lb < (uint32_t)BK_BUCKETS) {
    if constexpr (W == 1) {
      uint64_t k0, k1, k2, k3;
      lds_v2(sk + lb * 4, k0, k1);
      lds_v2(sk + lb * 4 + 2, k2, k3);
      int m = k3 == key[0] ? 3 : -1;
      m = k2 == key[0] ? 2 : m; m = k1 == key[0] ? 1 : m; m = k0 == key[0] ? 0 : m;
      uint32_t is_new = 0;
      if (m < 0) {
        int e = k3 == EMPTY_WORD ? 3 : -1;
        e = k2 == EMPTY_WORD ? 2 : e; e = k1 == EMPTY_WORD ? 1 : e; e = k0 == EMPTY_WORD ? 0 : e;
        if (e < 0) { ++lb; continue; }
        const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long *>(&sk[lb * 4 + e]), (unsigned long long)EMPTY_WORD,
                                       (unsigned long long)key[0]);
        if (old != EMPTY_WORD && old != key[0]) continue;        // another key took the slot: scan the same bucket again
        is_new = old == EMPTY_WORD ? 1u : 0u;
        m = e;
      }
      tile_value<MODE>(sv, lb * 4 + m, v, is_new, oldbits);
      return is_new;
    } else {
      uint64_t a0, a1, b0, b1;
      lds_v2(sk + lb * 4, a0, a1);
      lds_v2(sk + lb * 4 + 2, b0, b1);
      int m = (b0 == key[0] && b1 == key[W - 1]) ? 1 : -1;
      m = (a0 == key[0] && a1 == key[W - 1]) ? 0 : m;
      uint32_t is_new = 0;
      if (m < 0) {
        int e = (b0 == EMPTY_WORD && b1 == EMPTY_WORD) ? 1 : -1;
        e = (a0 == EMPTY_WORD && a1 == EMPTY_WORD) ? 0 : e;
        if (e < 0) { ++lb; continue; }
        uint64_t o0, o1;
        cas128_shared(sk + lb * 4 + e * 2, EMPTY_WORD, EMPTY_WORD, key[0], key[W - 1], o0, o1);
        const bool was_empty = o0 == EMPTY_WORD && o1 == EMPTY_WORD;
        if (!was_empty && !(o0 == key[0] && o1 == key[W - 1])) continue;
        is_new = was_empty ? 1u : 0u;
        m = e;
      }
      tile_value<MODE>(sv, lb * 2 + m, v, is_new, oldbits);
      return is_new;
    }
  }
  return 2u;
}

template <int W> constexpr size_t tile_smem_bytes() { return (size_t)BkGeom<W>::SLOTS * (8 * W + 4); }

// stage / write back one tile.  g: the tile's slots in the table
template <int W>
__device__ __forceinline__ void tile_stage(uint64_t *sk, uint32_t *sv, const uint4 *g, bool load) {
  constexpr int SLOTS = BkGeom<W>::SLOTS;
  if (!load) {
    for (int i = threadIdx.x; i < SLOTS; i += TI_THREADS) {
#pragma unroll
      for (int q = 0; q < W; ++q) sk[i * W + q] = EMPTY_WORD;
      sv[i] = 0u;
    }
  } else {
    for (int i = threadIdx.x; i < SLOTS; i += TI_THREADS) {
      uint64_t w[2 * W];
      LdCg<2 * W>::ld(g + (size_t)i * W, w);
#pragma unroll
      for (int q = 0; q < W; ++q) sk[i * W + q] = w[q];
      sv[i] = (uint32_t)w[W];
    }
  }
}
template <int W>
__device__ __forceinline__ void tile_writeback(const uint64_t *sk, const uint32_t *sv, uint4 *g) {
  constexpr int SLOTS = BkGeom<W>::SLOTS;
  for (int i = threadIdx.x; i < SLOTS; i += TI_THREADS) {
    if (W == 1) {
      const uint64_t k = sk[i];
      g[i] = make_uint4((uint32_t)k, (uint32_t)(k >> 32), sv[i], 0u);
    } else {
      const ulonglong2 k = *reinterpret_cast<const ulonglong2 *>(sk + 2 * i);
      g[(size_t)i * W] = make_uint4((uint32_t)k.x, (uint32_t)(k.x >> 32), (uint32_t)k.y, (uint32_t)(k.y >> 32));
      g[(size_t)i * W + W - 1] = make_uint4(sv[i], 0u, 0u, 0u);
    }
  }
}

template <int W>
__global__ void __launch_bounds__(TI_THREADS, 2) tile_insert_kernel(const TileArgs a) {
  constexpr int SLOTS = BkGeom<W>::SLOTS;
  constexpr int TI_MLP = TiMlp<W>::value;
  extern __shared__ __align__(16) unsigned char ti_smem[];
  uint64_t *sk = reinterpret_cast<uint64_t *>(ti_smem);
  uint32_t *sv = reinterpret_cast<uint32_t *>(sk + (size_t)SLOTS * W);
  const uint64_t n_tiles = a.t.capacity / SLOTS;
  const int lb_shift = 32 - a.lbk;                                 // bucket index = tag >> lb_shift (lbk <= 32)
  const uint32_t rpt = 1u << a.rpt_log2;
  uint32_t fresh = 0;

  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    uint4 *g = reinterpret_cast<uint4 *>(a.t.slots) + tile * (uint64_t)SLOTS * W;
    if (a.fresh == 1) {                                            // nothing to add to an empty, initialised tile
      bool any = false;
      if (a.slab) {
        for (uint32_t q = 0; q < rpt; ++q) {
          const uint64_t r = tile * rpt + q;
          any = any || a.ends[(size_t)r * BK_CSTRIDE] > r * a.slab;
        }
      } else any = a.ends[tile] > (tile ? a.ends[tile - 1] : 0ULL);
      if (!any) continue;                                          // block-uniform
    }
    tile_stage<W>(sk, sv, g, a.fresh == 0);
    __syncthreads();

    for (uint32_t q = 0; q < rpt; ++q) {
      uint32_t begin, end;                                         // record positions are 32-bit (the host keeps them < 2^32)
      if (a.slab) {
        const uint64_t r = tile * rpt + q;
        begin = (uint32_t)(r * a.slab);
        end = (uint32_t)min(a.ends[(size_t)r * BK_CSTRIDE], (unsigned long long)(r + 1) * a.slab);
      } else {
        begin = tile ? (uint32_t)a.ends[tile - 1] : 0u;
        end = (uint32_t)a.ends[tile];
      }
      // TI_MLP records per thread and step, all loads issued before the first probe (memory-level parallelism: the probes
      // are short, the loads are DRAM latency)
      for (uint64_t i0 = (uint64_t)begin + threadIdx.x; i0 < end; i0 += (uint64_t)TI_THREADS * TI_MLP) {
        uint4 rec[TI_MLP][W];
#pragma unroll
        for (int j = 0; j < TI_MLP; ++j) {
          const uint64_t i = i0 + (uint64_t)j * TI_THREADS;
          if (i < end) {
#pragma unroll
            for (int w = 0; w < W; ++w) rec[j][w] = __ldg(a.recs + i * W + w);
          }
        }
#pragma unroll
        for (int j = 0; j < TI_MLP; ++j) {
          const uint64_t i = i0 + (uint64_t)j * TI_THREADS;
          if (i < end) {
            uint64_t key[W];
            uint32_t tag;
            key[0] = (uint64_t)rec[j][0].x | ((uint64_t)rec[j][0].y << 32);
            if (W == 1) tag = rec[j][0].z;
            else { key[W - 1] = (uint64_t)rec[j][0].z | ((uint64_t)rec[j][0].w << 32); tag = rec[j][W - 1].x; }
            const uint32_t lb = (tag >> lb_shift) & (uint32_t)(BK_BUCKETS - 1);
            const uint32_t res = tile_put<W, TP_ADD>(sk, sv, key, lb, a.dflt);
            if (res == 2u) a.defer_idx[atomicAdd(a.n_new + 1, 1ULL)] = (uint32_t)i;
            else fresh += res;
          }
        }
      }
    }
    __syncthreads();
    tile_writeback<W>(sk, sv, g);
    __syncthreads();
  }
  fresh = warp_sum(fresh);
  if ((threadIdx.x & 31) == 0 && fresh) atomicAdd(a.n_new, (unsigned long long)fresh);
}

// The order-dependent reducers (FIRST = the reference's DiscardReducer / RH / LP insert, LAST = ReplaceReducer,
// KH/robinhood_offset_hashmap_ptr.hpp:85-97) on the same tiles.  The level-1 records carry their key's batch position
// (l1_scatter_kernel<W, true>).  Pass A claims the keys as above and races the positions in the VALUE word of the staged slot:
// FIRST: empty slots start at 0xffffffff, atomicMin(position); keys that were in the table before the batch (bitmap `old`)
// keep their value.  LAST: every slot starts at 0, atomicMax(position + 1); a key of the table that no record touched gets its
// value back from the table tile.  Pass B (per slot, after a block barrier) replaces the winning position by its value
// vals[position] (or dflt).  The write-back seals the occupied slots under FIRST (aux = AUX_SEALED_FIRST, as seal_kernel does)
// so that the general kernel, which takes the deferred keys afterwards, never overrides them.
template <int W> constexpr size_t tile_ord_smem_bytes() { return tile_smem_bytes<W>() + (size_t)BkGeom<W>::SLOTS / 8; }

template <int W, bool LAST>
__global__ void __launch_bounds__(TI_THREADS, 2) tile_insert_ord_kernel(const TileArgs a) {
  constexpr int SLOTS = BkGeom<W>::SLOTS;
  constexpr int TI_MLP = TiMlp<W>::value;
  static_assert(SLOTS % TI_THREADS == 0, "a warp stages 32 consecutive slots: one word of the bitmap");
  extern __shared__ __align__(16) unsigned char ti_smem[];
  uint64_t *sk = reinterpret_cast<uint64_t *>(ti_smem);
  uint32_t *sv = reinterpret_cast<uint32_t *>(sk + (size_t)SLOTS * W);
  uint32_t *old = sv + SLOTS;                                      // [SLOTS / 32] FIRST: the slot held a key before the batch
  const uint64_t n_tiles = a.t.capacity / SLOTS;
  const int lb_shift = 32 - a.lbk;
  const uint32_t rpt = 1u << a.rpt_log2;
  const bool load = a.fresh == 0;
  const uint32_t none = LAST ? 0u : 0xffffffffu;
  uint32_t fresh = 0;

  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    uint4 *g = reinterpret_cast<uint4 *>(a.t.slots) + tile * (uint64_t)SLOTS * W;
    if (a.fresh == 1) {
      bool any = false;
      if (a.slab) {
        for (uint32_t q = 0; q < rpt; ++q) {
          const uint64_t r = tile * rpt + q;
          any = any || a.ends[(size_t)r * BK_CSTRIDE] > r * a.slab;
        }
      } else any = a.ends[tile] > (tile ? a.ends[tile - 1] : 0ULL);
      if (!any) continue;
    }
    for (int i = threadIdx.x; i < SLOTS; i += TI_THREADS) {        // stage
      uint64_t w[2 * W];
      bool occ = false;
      if (load) {
        LdCg<2 * W>::ld(g + (size_t)i * W, w);
        occ = !key_is_empty<W>(w);
      } else {
#pragma unroll
        for (int q = 0; q < W; ++q) w[q] = EMPTY_WORD;
      }
#pragma unroll
      for (int q = 0; q < W; ++q) sk[i * W + q] = w[q];
      sv[i] = (!LAST && occ) ? (uint32_t)w[W] : none;
      if (!LAST) {
        const uint32_t m = __ballot_sync(FULL_MASK, occ);
        if ((threadIdx.x & 31) == 0) old[i >> 5] = m;
      }
    }
    __syncthreads();

    for (uint32_t q = 0; q < rpt; ++q) {                           // pass A
      uint32_t begin, end;
      if (a.slab) {
        const uint64_t r = tile * rpt + q;
        begin = (uint32_t)(r * a.slab);
        end = (uint32_t)min(a.ends[(size_t)r * BK_CSTRIDE], (unsigned long long)(r + 1) * a.slab);
      } else {
        begin = tile ? (uint32_t)a.ends[tile - 1] : 0u;
        end = (uint32_t)a.ends[tile];
      }
      for (uint64_t i0 = (uint64_t)begin + threadIdx.x; i0 < end; i0 += (uint64_t)TI_THREADS * TI_MLP) {
        uint4 rec[TI_MLP][W];
#pragma unroll
        for (int j = 0; j < TI_MLP; ++j) {
          const uint64_t i = i0 + (uint64_t)j * TI_THREADS;
          if (i < end) {
#pragma unroll
            for (int w = 0; w < W; ++w) rec[j][w] = __ldg(a.recs + i * W + w);
          }
        }
#pragma unroll
        for (int j = 0; j < TI_MLP; ++j) {
          const uint64_t i = i0 + (uint64_t)j * TI_THREADS;
          if (i < end) {
            uint64_t key[W];
            uint32_t tag, pos;
            key[0] = (uint64_t)rec[j][0].x | ((uint64_t)rec[j][0].y << 32);
            if (W == 1) { tag = rec[j][0].z; pos = rec[j][0].w; }
            else { key[W - 1] = (uint64_t)rec[j][0].z | ((uint64_t)rec[j][0].w << 32); tag = rec[j][W - 1].x; pos = rec[j][W - 1].y; }
            const uint32_t lb = (tag >> lb_shift) & (uint32_t)(BK_BUCKETS - 1);
            const uint32_t res = LAST ? tile_put<W, TP_MAX>(sk, sv, key, lb, pos + 1u)
                                      : tile_put<W, TP_MIN>(sk, sv, key, lb, pos, load ? old : nullptr);
            if (res == 2u) a.defer_idx[atomicAdd(a.n_new + 1, 1ULL)] = (uint32_t)i;
            else fresh += res;
          }
        }
      }
    }
    __syncthreads();

    for (int i = threadIdx.x; i < SLOTS; i += TI_THREADS) {        // pass B + write back
      uint64_t k[W];
#pragma unroll
      for (int q = 0; q < W; ++q) k[q] = sk[i * W + q];
      const bool occ = !key_is_empty<W>(k);
      uint32_t v = sv[i];
      if (!occ) v = 0u;
      else if (LAST) {
        if (v) v = a.vals ? __ldg(a.vals + (v - 1u)) : a.dflt;
        else v = W == 1 ? __ldcg(reinterpret_cast<const uint32_t *>(g + i) + 2) : __ldcg(reinterpret_cast<const uint32_t *>(g + (size_t)i * W + W - 1));
      } else if (!(load && ((old[i >> 5] >> (i & 31)) & 1u))) {
        v = a.vals ? __ldg(a.vals + v) : a.dflt;
      }
      const uint32_t aux = (!LAST && occ) ? AUX_SEALED_FIRST : 0u;
      if (W == 1) g[i] = make_uint4((uint32_t)k[0], (uint32_t)(k[0] >> 32), v, aux);
      else {
        g[(size_t)i * W] = make_uint4((uint32_t)k[0], (uint32_t)(k[0] >> 32), (uint32_t)k[W - 1], (uint32_t)(k[W - 1] >> 32));
        g[(size_t)i * W + W - 1] = make_uint4(v, aux, 0u, 0u);
      }
    }
    __syncthreads();
  }
  fresh = warp_sum(fresh);
  if ((threadIdx.x & 31) == 0 && fresh) atomicAdd(a.n_new, (unsigned long long)fresh);
}

// (key, position) of the deferred records of an ordered batch
template <int W>
__global__ void gather_rec_pos_kernel(const uint4 *recs, const uint32_t *idx, uint64_t n, uint32_t *pos) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 r = recs[(size_t)idx[i] * W + W - 1];
  pos[i] = W == 1 ? r.w : r.y;
}
template <int W>
__global__ void gather_rec_pairs_kernel(const uint4 *recs, const uint32_t *idx, const uint32_t *pos, const uint32_t *vals, uint64_t n,
                                        uint64_t *out_keys, uint32_t *out_vals) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 r = recs[(size_t)idx[i] * W];
  out_keys[i * W] = (uint64_t)r.x | ((uint64_t)r.y << 32);
  if (W == 2) out_keys[i * W + 1] = (uint64_t)r.z | ((uint64_t)r.w << 32);
  if (vals) out_vals[i] = vals[pos[i]];
}

// ---------------------------------------------------------------------------------------------------------------------
// region insert: a thread-block CLUSTER per level-1 region, for tables with more tiles than regions
// ---------------------------------------------------------------------------------------------------------------------
// When a region spans T = 2^tb tiles (1 <= tb <= 4), the level-2 pass -- another read and write of every record -- can
// be saved: a cluster of T CTAs takes the region, CTA c stages tile (region << tb) + c in ITS shared memory, all CTAs
// stream the region's records (interleaved chunks) and every record is inserted into the tile its tag names, local or
// in a neighbour SM's shared memory (distributed shared memory: cluster.map_shared_rank + generic loads / atomics).
// Probe rule, deferral and write-back are the tile insert's; the cluster barrier separates stage / insert / write-back.

// 16 bytes in ONE load through a generic pointer (a key of two words must not be seen half written by a neighbour's 128-bit CAS)
__device__ __forceinline__ void ldg_v2(const uint64_t *p, uint64_t &a, uint64_t &b) {
  asm volatile("ld.volatile.v2.u64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
}
// tile_put on GENERIC pointers (the tile may live in another CTA's shared memory)
template <int W>
__device__ __forceinline__ uint32_t tile_put_cluster(uint64_t *sk, uint32_t *sv, const uint64_t (&key)[W], uint32_t lb, uint32_t v) {
  if (key_is_empty<W>(key)) return 2u;
  while (lb < (uint32_t)BK_BUCKETS) {
    if constexpr (W == 1) {
      uint64_t k0, k1, k2, k3;
      ldg_v2(sk + lb * 4, k0, k1);
      ldg_v2(sk + lb * 4 + 2, k2, k3);
      int m = k3 == key[0] ? 3 : -1;
      m = k2 == key[0] ? 2 : m; m = k1 == key[0] ? 1 : m; m = k0 == key[0] ? 0 : m;
      uint32_t is_new = 0;
      if (m < 0) {
        int e = k3 == EMPTY_WORD ? 3 : -1;
        e = k2 == EMPTY_WORD ? 2 : e; e = k1 == EMPTY_WORD ? 1 : e; e = k0 == EMPTY_WORD ? 0 : e;
        if (e < 0) { ++lb; continue; }
        const uint64_t old = atomicCAS(reinterpret_cast<unsigned long long *>(&sk[lb * 4 + e]), (unsigned long long)EMPTY_WORD,
                                       (unsigned long long)key[0]);
        if (old != EMPTY_WORD && old != key[0]) continue;
        is_new = old == EMPTY_WORD ? 1u : 0u;
        m = e;
      }
      atomicAdd(&sv[lb * 4 + m], v);
      return is_new;
    } else {
      uint64_t a0, a1, b0, b1;
      ldg_v2(sk + lb * 4, a0, a1);
      ldg_v2(sk + lb * 4 + 2, b0, b1);
      int m = (b0 == key[0] && b1 == key[W - 1]) ? 1 : -1;
      m = (a0 == key[0] && a1 == key[W - 1]) ? 0 : m;
      uint32_t is_new = 0;
      if (m < 0) {
        int e = (b0 == EMPTY_WORD && b1 == EMPTY_WORD) ? 1 : -1;
        e = (a0 == EMPTY_WORD && a1 == EMPTY_WORD) ? 0 : e;
        if (e < 0) { ++lb; continue; }
        uint64_t o0, o1;
        uint64_t *p = sk + lb * 4 + e * 2;
        asm volatile("{\n\t.reg .b128 c, n, o;\n\tmov.b128 c, {%3, %4};\n\tmov.b128 n, {%5, %6};\n\t"
                     "atom.relaxed.cluster.cas.b128 o, [%2], c, n;\n\tmov.b128 {%0, %1}, o;\n\t}"
                     : "=l"(o0), "=l"(o1) : "l"(p), "l"(EMPTY_WORD), "l"(EMPTY_WORD), "l"(key[0]), "l"(key[W - 1]) : "memory");
        const bool was_empty = o0 == EMPTY_WORD && o1 == EMPTY_WORD;
        if (!was_empty && !(o0 == key[0] && o1 == key[W - 1])) continue;
        is_new = was_empty ? 1u : 0u;
        m = e;
      }
      atomicAdd(&sv[lb * 2 + m], v);
      return is_new;
    }
  }
  return 2u;
}

template <int W>
__global__ void __launch_bounds__(TI_THREADS, 2) region_insert_kernel(const TileArgs a, const int tb) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  constexpr int SLOTS = BkGeom<W>::SLOTS;
  constexpr int TI_MLP = TiMlp<W>::value;
  extern __shared__ __align__(16) unsigned char ti_smem[];
  uint64_t *sk = reinterpret_cast<uint64_t *>(ti_smem);
  uint32_t *sv = reinterpret_cast<uint32_t *>(sk + (size_t)SLOTS * W);
  const uint32_t T = 1u << tb;                                     // == cluster.num_blocks()
  const uint32_t crank = cluster.block_rank();
  const uint64_t n_regions = (a.t.capacity / SLOTS) >> tb;
  const uint64_t cid = blockIdx.x >> tb, ncl = gridDim.x >> tb;
  const int lb_shift = 32 - a.lbk;                                 // table bucket = tag >> lb_shift
  constexpr uint32_t CH = TI_THREADS * TI_MLP;                     // records per CTA and step
  uint32_t fresh = 0;

  for (uint64_t region = cid; region < n_regions; region += ncl) {
    const uint32_t begin = (uint32_t)(region * a.slab);
    const uint32_t end = (uint32_t)min(a.ends[(size_t)region * BK_CSTRIDE], (unsigned long long)(region + 1) * a.slab);
    if (a.fresh == 1 && begin == end) continue;                    // cluster-uniform: nothing to add to empty, initialised tiles
    const uint64_t tile = (region << tb) + crank;
    uint4 *g = reinterpret_cast<uint4 *>(a.t.slots) + tile * (uint64_t)SLOTS * W;
    tile_stage<W>(sk, sv, g, a.fresh == 0);
    cluster.sync();
    for (uint64_t i0 = (uint64_t)begin + (uint64_t)crank * CH + threadIdx.x; i0 < end; i0 += (uint64_t)T * CH) {
      uint4 rec[TI_MLP][W];
#pragma unroll
      for (int j = 0; j < TI_MLP; ++j) {
        const uint64_t i = i0 + (uint64_t)j * TI_THREADS;
        if (i < end) {
#pragma unroll
          for (int w = 0; w < W; ++w) rec[j][w] = __ldg(a.recs + i * W + w);
        }
      }
#pragma unroll
      for (int j = 0; j < TI_MLP; ++j) {
        const uint64_t i = i0 + (uint64_t)j * TI_THREADS;
        if (i < end) {
          uint64_t key[W];
          uint32_t tag;
          key[0] = (uint64_t)rec[j][0].x | ((uint64_t)rec[j][0].y << 32);
          if (W == 1) tag = rec[j][0].z;
          else { key[W - 1] = (uint64_t)rec[j][0].z | ((uint64_t)rec[j][0].w << 32); tag = rec[j][W - 1].x; }
          const uint32_t gb = tag >> lb_shift;                     // bucket of the table
          const uint32_t owner = (gb >> BK_BUCKET_BITS) & (T - 1u);
          uint64_t *rk = cluster.map_shared_rank(sk, owner);
          uint32_t *rv = cluster.map_shared_rank(sv, owner);
          const uint32_t res = tile_put_cluster<W>(rk, rv, key, gb & (uint32_t)(BK_BUCKETS - 1), a.dflt);
          if (res == 2u) a.defer_idx[atomicAdd(a.n_new + 1, 1ULL)] = (uint32_t)i;
          else fresh += res;
        }
      }
    }
    cluster.sync();
    tile_writeback<W>(sk, sv, g);
    __syncthreads();
  }
  fresh = warp_sum(fresh);
  if ((threadIdx.x & 31) == 0 && fresh) atomicAdd(a.n_new, (unsigned long long)fresh);
}


// =====================================================================================================================
// rebuild: old table -> new table as one streaming pass (reserve / rehash / shrink-after-erase / sticky cleaning)
// KH/robinhood_offset_hashmap_ptr.hpp:907-981 (rehash), :1122-1331 (copy_upsize), :1055-1111 (copy_downsize)
// =====================================================================================================================
struct RebuildArgs {
  TableView src, dst;                    // dst.capacity >= BkGeom<W>::SLOTS; both powers of two
  int hash_kind;
  HashParams hp;
  int lbk;                               // log2(buckets of dst)
  unsigned long long *n_placed;          // [0] += entries placed, [1] += entries deferred
  uint64_t *defer_keys;                  // entries whose home bucket is not in the tile fed by their old slot, or whose chain
  uint32_t *defer_vals;                  // leaves it: (key, value) for the global insert kernel
  unsigned long long defer_cap;
};

// The old slots are 30-60 % occupied, so a warp that hashed and placed "its" slots would run the expensive part (hash, probe,
// claim: ~150 instructions) with most lanes idle -- and this kernel is issue-bound (ncu r02e: sm throughput 65 %, dram 10 %).
// Instead every warp packs the occupied entries of the 32 * TI_MLP slots it loaded into a small stage in shared memory
// (ballot + prefix) and works through them with all lanes busy.
constexpr int RB_STAGE = 64;                                      // entries per warp and pass (two loads per lane)
template <int W> constexpr size_t rebuild_smem_bytes() { return tile_smem_bytes<W>() + (size_t)(TI_THREADS / 32) * RB_STAGE * (8 * W + 4); }

template <int W>
__global__ void __launch_bounds__(TI_THREADS, 2) tile_rebuild_kernel(const RebuildArgs a) {
  constexpr int SLOTS = BkGeom<W>::SLOTS;
  constexpr int TI_MLP = TiMlp<W>::value;
  constexpr int JP = RB_STAGE / 32;                                // loads per lane and pass
  static_assert(TI_MLP % JP == 0, "a pass stages whole loads");
  extern __shared__ __align__(16) unsigned char ti_smem[];
  uint64_t *sk = reinterpret_cast<uint64_t *>(ti_smem);
  uint32_t *sv = reinterpret_cast<uint32_t *>(sk + (size_t)SLOTS * W);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint64_t *stk = reinterpret_cast<uint64_t *>(sv + SLOTS) + (size_t)warp * RB_STAGE * W;          // this warp's stage: keys
  uint32_t *stv = reinterpret_cast<uint32_t *>(reinterpret_cast<uint64_t *>(sv + SLOTS) + (size_t)(TI_THREADS / 32) * RB_STAGE * W) +
                  warp * RB_STAGE;                                                                  // ... and values
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint64_t n_tiles = a.dst.capacity / SLOTS;
  // old slots feeding new tile t: [t * per, (t + 1) * per), per = SLOTS * src.capacity / dst.capacity
  const uint64_t per_num = (uint64_t)SLOTS * a.src.capacity;       // < 2^64: capacities <= 2^40
  const uint64_t per = per_num / a.dst.capacity;                   // 0 when the table grows by more than SLOTS x
  const uint64_t tiles_per_slot = per ? 1 : a.dst.capacity / per_num;
  const int bshift = 64 - a.lbk;
  uint32_t placed = 0;
  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    uint4 *g = reinterpret_cast<uint4 *>(a.dst.slots) + tile * (uint64_t)SLOTS * W;
    tile_stage<W>(sk, sv, g, false);
    __syncthreads();
    uint64_t s0, s1;
    if (per) { s0 = tile * per; s1 = s0 + per; }
    else {                                                         // several new tiles per old slot: the first of them takes it
      s0 = tile / tiles_per_slot; s1 = (tile % tiles_per_slot == 0) ? s0 + 1 : s0;
    }
    const uint4 *old = reinterpret_cast<const uint4 *>(a.src.slots);
    // warp-uniform trip count: all lanes of a warp share i0 - lane
    for (uint64_t i0 = s0 + threadIdx.x; i0 - lane < s1; i0 += (uint64_t)TI_THREADS * TI_MLP) {
      uint64_t w[TI_MLP][2 * W];
#pragma unroll
      for (int j = 0; j < TI_MLP; ++j) {
        const uint64_t i = i0 + (uint64_t)j * TI_THREADS;
#pragma unroll
        for (int q = 0; q < 2 * W; ++q) w[j][q] = EMPTY_WORD;
        if (i < s1) LdCg<2 * W>::ld(old + i * W, w[j]);
      }
#pragma unroll
      for (int j0 = 0; j0 < TI_MLP; j0 += JP) {
        uint32_t total = 0;
#pragma unroll
        for (int jj = 0; jj < JP; ++jj) {
          const bool occ = !key_is_empty<W>(w[j0 + jj]);
          const uint32_t m = __ballot_sync(FULL_MASK, occ);
          if (occ) {
            const uint32_t pos = total + __popc(m & lt_mask);
#pragma unroll
            for (int q = 0; q < W; ++q) stk[pos * W + q] = w[j0 + jj][q];
            stv[pos] = (uint32_t)w[j0 + jj][W];
          }
          total += __popc(m);
        }
        __syncwarp();
        for (uint32_t e = lane; e < total; e += 32) {
          uint64_t key[W];
#pragma unroll
          for (int q = 0; q < W; ++q) key[q] = stk[e * W + q];
          const uint32_t val = stv[e];
          const uint64_t b = top_bits(spread_hash(hash_runtime<W>(a.hash_kind, key, a.hp)), bshift);
          uint32_t res = 2u;
          if ((b >> BK_BUCKET_BITS) == tile) res = tile_put<W, TP_SET>(sk, sv, key, (uint32_t)b & (uint32_t)(BK_BUCKETS - 1), val);
          if (res == 2u) {
            const unsigned long long at = atomicAdd(a.n_placed + 1, 1ULL);
            if (at < a.defer_cap) {
#pragma unroll
              for (int q = 0; q < W; ++q) a.defer_keys[at * W + q] = key[q];
              a.defer_vals[at] = val;
            }
          } else placed += res;
        }
        __syncwarp();
      }
    }
    __syncthreads();
    tile_writeback<W>(sk, sv, g);
    __syncthreads();
  }
  placed = warp_sum(placed);
  if ((threadIdx.x & 31) == 0 && placed) atomicAdd(a.n_placed, (unsigned long long)placed);
}

}  // namespace kh
