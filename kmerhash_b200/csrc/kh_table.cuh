// kh_table.cuh -- the open-addressing k-mer table in HBM and its probe kernels.
//
// What it replaces (semantics, not structure): the reference's three CPU tables
//   hashmap_robinhood_offsets_reduction  KH/robinhood_offset_hashmap_ptr.hpp:196-202
//     insert_with_hint :1522-1649, find_pos_with_hint :1339-1395, erase_and_compact :3504-3556
//   hashmap_robinhood_doubling           KH/hashmap_robinhood.hpp:126 (insert :522-624, find_pos :1058-1095)
//   hashmap_linearprobe_doubling         KH/hashmap_linearprobe.hpp:98 (insert :430-517, erase :911-978)
// Those walk one key at a time and hide two dependent cache misses per key behind software
// prefetch.  Here the table is an array of 16-byte (k <= 32) or 32-byte (k <= 64) slots
//   { key[W] : u64 x W, val : u32, aux : u32 (, pad : u64 when W == 2) }
// grouped into buckets of 32/64/128 bytes.  A bucket is read with ONE coalesced access by a
// group of G lanes (each lane one 16- or 32-byte vector load), matched with warp ballots,
// claimed with a 64-bit (W=1) or 128-bit (W=2) atomicCAS on the key and counted with a
// fire-and-forget atomicAdd (RED) on val.  Buckets are probed linearly.  Thousands of keys are
// in flight per SM, which is what hides HBM latency here.
//
// Concurrency rules (what makes the batch kernels safe):
//   * insert phase: an occupied slot never changes its key, every thread claims the FIRST
//     empty slot in probe order and re-scans when it loses a CAS  =>  no key is stored twice.
//   * erase is tombstone-free: the slot becomes EMPTY again and is reusable at once.  When a
//     key leaves a bucket that was full, the slot's aux keeps a STICKY bit ("a probe chain
//     may run through this bucket"); lookups continue past a bucket iff it is full or has a
//     sticky slot.  A bulk rebuild (rehash) clears all sticky bits.
//   * insert, lookup and erase are separate batch phases (as in every reference driver).
//   * all table reads go through L2 (ld.global.cg) so they see other SMs' atomics.
//
// Order-dependent reducers (FIRST = DiscardReducer / RH / LP insert, LAST = ReplaceReducer,
// KH/robinhood_offset_hashmap_ptr.hpp:85-97): phase A claims keys and races the batch
// position into aux with atomicMax, phase B lets the winning position store its value.
#pragma once
#include <cstdint>
#include "kh_hash.cuh"

namespace kh {

constexpr uint64_t EMPTY_WORD = ~0ULL;
constexpr uint32_t AUX_STICKY = 0x80000000u;
constexpr uint32_t AUX_IDX_MASK = 0x7fffffffu;
constexpr uint32_t AUX_SEALED_FIRST = 0x7fffffffu;   // pre-existing key under FIRST: never loses
constexpr uint32_t FULL_MASK = 0xffffffffu;

// Home bucket = the TOP log2(n_buckets) bits of spread(hash).  The multiplicative spread makes the top
// bits depend on every hash bit (32-bit functors and the identity functor leave the top bits of the raw
// value empty).  Taking the top bits -- not `hash & mask` as the reference does -- makes the order of
// the home buckets independent of the capacity: "region" r of P (top log2 P bits) is one contiguous
// 1/P of the table at ANY capacity, which is what the L2-windowed bulk insert sorts by, and a table
// dump in slot order is already sorted by region for the rebuilt table.
__host__ __device__ __forceinline__ uint64_t spread_hash(uint64_t h) { return h * 0x9E3779B97F4A7C15ULL; }
__host__ __device__ __forceinline__ uint64_t top_bits(uint64_t hm, int shift) { return shift >= 64 ? 0 : (hm >> shift); }
__host__ __device__ __forceinline__ int log2_u64(uint64_t x) {   // x is a power of two
#ifdef __CUDA_ARCH__
  return 63 - __clzll((long long)x);
#else
  int l = 0; while ((1ULL << l) < x) ++l; return l;
#endif
}

enum : int { OP_FIND = 0, OP_COUNT = 1, OP_ERASE = 2, OP_UPDATE = 3, OP_FINALIZE = 4 };
enum : int { INS_COUNT = 0, INS_FIRST = 1, INS_LAST = 2 };

struct TableView {
  unsigned char *slots;   // capacity * slot_bytes, 256-byte aligned
  uint64_t capacity;      // number of slots, power of two
  uint32_t *side;         // side slot for the all-ones key: [0] present, [1] val, [2] aux idx
};

template <int W> struct SlotBytes { static constexpr int value = 16 * W; };

// ---- vector loads that bypass L1 ---------------------------------------------------------
template <int NW> struct LdCg;
template <> struct LdCg<2> {
  static __device__ __forceinline__ void ld(const void *p, uint64_t (&w)[2]) {
    asm volatile("ld.global.cg.v2.u64 {%0,%1}, [%2];" : "=l"(w[0]), "=l"(w[1]) : "l"(p) : "memory");
  }
};
template <> struct LdCg<4> {
  static __device__ __forceinline__ void ld(const void *p, uint64_t (&w)[4]) {
    asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3]) : "l"(p) : "memory");
  }
};

template <> struct LdCg<8> {   // a whole 64-byte bucket by one thread: two 32-byte loads in flight
  static __device__ __forceinline__ void ld(const void *p, uint64_t (&w)[8]) {
    asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(w[0]), "=l"(w[1]), "=l"(w[2]), "=l"(w[3]) : "l"(p) : "memory");
    asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(w[4]), "=l"(w[5]), "=l"(w[6]), "=l"(w[7]) : "l"((const char *)p + 32) : "memory");
  }
};

__device__ __forceinline__ uint32_t ldcg_u32(const uint32_t *p) {
  uint32_t v; asm volatile("ld.global.cg.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}

// ---- key claim ----------------------------------------------------------------------------
// returns true when the slot now holds `key` because of us (claimed) ; `same` = it already held key
template <int W>
__device__ __forceinline__ int cas_key(void *addr, const uint64_t (&expect)[W], const uint64_t (&desired)[W],
                                       uint64_t (&old)[W]) {
  if (W == 1) {
    old[0] = atomicCAS((unsigned long long *)addr, (unsigned long long)expect[0], (unsigned long long)desired[0]);
  } else {
    uint64_t olo, ohi;
    asm volatile("{\n\t.reg .b128 c, n, o;\n\tmov.b128 c, {%3, %4};\n\tmov.b128 n, {%5, %6};\n\t"
                 "atom.relaxed.gpu.global.cas.b128 o, [%2], c, n;\n\tmov.b128 {%0, %1}, o;\n\t}"
                 : "=l"(olo), "=l"(ohi)
                 : "l"(addr), "l"(expect[0]), "l"(expect[W - 1]), "l"(desired[0]), "l"(desired[W - 1])
                 : "memory");
    old[0] = olo; old[W - 1] = ohi;
  }
  bool was_expected = true;
#pragma unroll
  for (int i = 0; i < W; ++i) was_expected = was_expected && (old[i] == expect[i]);
  return was_expected;
}

template <int W> __device__ __forceinline__ bool key_eq(const uint64_t *a, const uint64_t (&b)[W]) {
  bool e = true;
#pragma unroll
  for (int i = 0; i < W; ++i) e = e && (a[i] == b[i]);
  return e;
}
template <int W> __device__ __forceinline__ bool key_is_empty(const uint64_t *a) {
  bool e = true;
#pragma unroll
  for (int i = 0; i < W; ++i) e = e && (a[i] == EMPTY_WORD);
  return e;
}

template <int W>
__device__ __forceinline__ void load_key(const uint64_t *keys, uint64_t i, uint64_t (&k)[W]) {
  if (W == 1) k[0] = __ldg(keys + i);
  else { ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2 *>(keys) + i); k[0] = v.x; k[W - 1] = v.y; }
}
template <int W>
__device__ __forceinline__ void shfl_key(const uint64_t (&k)[W], int src, uint64_t (&o)[W]) {
#pragma unroll
  for (int i = 0; i < W; ++i) o[i] = __shfl_sync(FULL_MASK, k[i], src);
}

template <int W>
__device__ __forceinline__ uint64_t hash_runtime(int kind, const uint64_t (&key)[W], const HashParams &hp) {
  // The default store / rank hash first, as a plain uniform branch: with only the switch in the way the compiler keeps
  // every key's hash in its own basic block and the unrolled per-thread batches (8 keys in the region scatter) lose their
  // instruction-level parallelism -- the scatter ran 1.45 ms with the switch alone, 1.21 ms with this line.
  if (kind == KH_HASH_MURMUR64AVX) return hash_key<KH_HASH_MURMUR64AVX, W>(key, hp);
  uint64_t h = 0;
  switch (kind) {
    case KH_HASH_IDEN:        h = hash_key<KH_HASH_IDEN, W>(key, hp); break;
    case KH_HASH_MURMUR64:    h = hash_key<KH_HASH_MURMUR64, W>(key, hp); break;
    case KH_HASH_MURMUR64AVX: h = hash_key<KH_HASH_MURMUR64AVX, W>(key, hp); break;
    case KH_HASH_MURMUR32:    h = hash_key<KH_HASH_MURMUR32, W>(key, hp); break;
    case KH_HASH_FARM:        h = hash_key<KH_HASH_FARM, W>(key, hp); break;
    case KH_HASH_FARM32:      h = hash_key<KH_HASH_FARM32, W>(key, hp); break;
    case KH_HASH_CLHASH:      h = hash_key<KH_HASH_CLHASH, W>(key, hp); break;
    case KH_HASH_CRC32C:      h = hash_key<KH_HASH_CRC32C, W>(key, hp); break;
    default: break;
  }
  return h;
}

// Geometry: W words per key, LB bytes per lane load (16, 32 or 64), G lanes per bucket.
//   slots per lane   S = LB / (16 W)        (W=1: 1, 2 or 4; W=2: 1 or 2, LB >= 32)
//   slots per bucket  = G * S,  bucket bytes = G * LB
template <int W, int LB, int G>
struct Geom {
  static constexpr int SLOT = 16 * W;
  static constexpr int S = LB / SLOT;
  static constexpr int NW = LB / 8;           // u64 words per lane load
  static constexpr int SW = SLOT / 8;         // u64 words per slot
  static constexpr int BUCKET_SLOTS = G * S;
  static_assert(S >= 1, "lane load smaller than a slot");
};

// What one lane learns from its LB bytes of a bucket
template <int W, int LB, int G>
struct LaneScan {
  int match;      // local slot index holding the key, or -1
  int empty;      // local index of the first empty slot, or -1
  bool sticky;
  uint64_t va;    // val|aux word of the matching slot (valid when match >= 0)
};

template <int W, int LB, int G>
__device__ __forceinline__ LaneScan<W, LB, G> scan_lane(const unsigned char *lane_ptr, const uint64_t (&key)[W]) {
  using Ge = Geom<W, LB, G>;
  uint64_t w[Ge::NW];
  LdCg<Ge::NW>::ld(lane_ptr, w);
  LaneScan<W, LB, G> r; r.match = -1; r.empty = -1; r.sticky = false; r.va = 0;
#pragma unroll
  for (int j = Ge::S - 1; j >= 0; --j) {
    const uint64_t *sk = &w[j * Ge::SW];
    const uint64_t va = w[j * Ge::SW + W];
    if (key_eq<W>(sk, key)) { r.match = j; r.va = va; }
    if (key_is_empty<W>(sk)) r.empty = j;
    r.sticky = r.sticky || (((uint32_t)(va >> 32)) & AUX_STICKY);
  }
  return r;
}

// group-wide vote helpers: G == 1 needs no warp traffic at all
template <int G>
__device__ __forceinline__ uint32_t group_ballot(bool pred, int gshift) {
  if (G == 1) return pred ? 1u : 0u;
  uint32_t b = __ballot_sync(FULL_MASK, pred);
  return (b >> gshift) & ((G == 32) ? FULL_MASK : ((1u << G) - 1u));
}
template <int G>
__device__ __forceinline__ bool warp_any(bool pred) {
  if (G == 1) return pred;            // divergent per-thread loop
  return __any_sync(FULL_MASK, pred);
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
  return v;
}

// =========================================================================================
// lookup-type operations: find / count / erase / update / finalize(ordered insert phase B)
// =========================================================================================
struct LookupArgs {
  TableView t;
  const uint64_t *keys;
  uint64_t n;
  int hash_kind;
  HashParams hp;
  // OP_FIND: out_u32 + nonexistent; OP_COUNT: out_u8; OP_UPDATE / OP_FINALIZE: vals (may be null -> dflt)
  uint32_t *out_u32;
  uint8_t *out_u8;
  const uint32_t *vals;
  uint32_t nonexistent;     // OP_FIND fill; OP_FINALIZE/UPDATE default value
  uint32_t ins_mode;        // OP_FINALIZE: INS_FIRST / INS_LAST
  uint64_t idx_base;        // OP_FINALIZE: batch position of keys[0] (encoded positions restart per chunk)
  unsigned long long *n_hit;   // found / erased / updated counter
};

template <int W, int LB, int G, int OP>
__global__ void __launch_bounds__(256) lookup_kernel(LookupArgs a) {
  using Ge = Geom<W, LB, G>;
  const int lane = threadIdx.x & 31;
  const int gl = lane & (G - 1);
  const int gbase = lane - gl;
  const uint64_t bmask = a.t.capacity / Ge::BUCKET_SLOTS - 1;
  const int bshift = 64 - log2_u64(bmask + 1);
  const uint64_t warps_total = (uint64_t)gridDim.x * (blockDim.x >> 5);
  const uint64_t warp_id = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  uint32_t hits = 0;

  for (uint64_t base = warp_id * 32; base < a.n; base += warps_total * 32) {
    const uint64_t i = base + lane;
    const bool valid = i < a.n;
    uint64_t key[W];
#pragma unroll
    for (int q = 0; q < W; ++q) key[q] = 0;
    if (valid) load_key<W>(a.keys, i, key);
    bool probe = valid;
    bool found = false;
    uint32_t fval = 0;

    // the all-ones key lives in the side slot
    if (valid && key_is_empty<W>(key)) {
      probe = false;
      if (OP == OP_ERASE) {
        if (atomicExch(&a.t.side[0], 0u) == 1u) { found = true; ++hits; a.t.side[1] = 0; a.t.side[2] = 0; }
      } else if (ldcg_u32(&a.t.side[0]) == 1u) {
        if (OP == OP_FIND || OP == OP_COUNT) { found = true; fval = ldcg_u32(&a.t.side[1]); }
        else if (OP == OP_UPDATE) { found = true; a.t.side[1] = a.vals ? a.vals[i] : a.nonexistent; }
        else if (OP == OP_FINALIZE) {
          const uint32_t pos = (uint32_t)(i - a.idx_base);
          const uint32_t enc = (a.ins_mode == INS_FIRST) ? (0x7ffffffeu - pos) : (pos + 1u);
          if (ldcg_u32(&a.t.side[2]) == enc) {
            a.t.side[1] = a.vals ? a.vals[i] : a.nonexistent;
            a.t.side[2] = (a.ins_mode == INS_FIRST) ? AUX_SEALED_FIRST : 0u;
          }
        }
      }
    }
    const uint64_t h = top_bits(spread_hash(hash_runtime<W>(a.hash_kind, key, a.hp)), bshift);   // home bucket

#pragma unroll 1
    for (int r = 0; r < G; ++r) {
      uint64_t k[W];
      uint64_t b;
      bool act;
      if (G == 1) {
#pragma unroll
        for (int q = 0; q < W; ++q) k[q] = key[q];
        b = h; act = probe;
      } else {
        shfl_key<W>(key, gbase + r, k);
        b = __shfl_sync(FULL_MASK, h, gbase + r);
        act = __shfl_sync(FULL_MASK, (int)probe, gbase + r) != 0;
      }
      uint64_t probes = 0;
      bool gfound = false;
      uint32_t gval = 0;
      while (warp_any<G>(act)) {
        LaneScan<W, LB, G> s; s.match = -1; s.empty = -1; s.sticky = false; s.va = 0;
        unsigned char *lane_ptr = a.t.slots + (b * G + gl) * (uint64_t)LB;
        if (act) s = scan_lane<W, LB, G>(lane_ptr, k);
        const uint32_t gm = group_ballot<G>(act && s.match >= 0, gbase);
        const uint32_t ge = group_ballot<G>(act && s.empty >= 0, gbase);
        const uint32_t gs = group_ballot<G>(act && s.sticky, gbase);
        if (OP == OP_FIND) {   // hand the value of the matching lane to the whole group
          uint32_t v = (uint32_t)s.va;
          if (G > 1) v = __shfl_sync(FULL_MASK, v, gm ? (gbase + __ffs(gm) - 1) : lane);
          if (act && gm) gval = v;
        }
        if (act) {
          if (gm) {
            if (s.match >= 0) {   // exactly one lane: the owner of the slot
              unsigned char *slot = lane_ptr + s.match * Ge::SLOT;
              uint32_t *pval = reinterpret_cast<uint32_t *>(slot + 8 * W);
              uint32_t *paux = pval + 1;
              const uint32_t aux = (uint32_t)(s.va >> 32);
              if (OP == OP_ERASE) {
                // sticky first (if the bucket is full now), then release the key
                uint32_t keep = aux & AUX_STICKY;
                if (ge == 0 && !keep) { keep = AUX_STICKY; atomicOr(paux, AUX_STICKY); __threadfence(); }
                uint64_t expect[W], empty[W], old[W];
#pragma unroll
                for (int q = 0; q < W; ++q) { expect[q] = k[q]; empty[q] = EMPTY_WORD; }
                if (cas_key<W>(slot, expect, empty, old)) {
                  // we erased it: reset val and the ordered-insert position, keep sticky
                  *reinterpret_cast<volatile uint64_t *>(pval) = (uint64_t)keep << 32;
                  ++hits;
                }
              } else if (OP == OP_UPDATE) {
                *reinterpret_cast<volatile uint32_t *>(pval) = a.vals ? a.vals[base + gbase + r] : a.nonexistent;
              } else if (OP == OP_FINALIZE) {
                const uint64_t src_i = base + gbase + r;
                const uint32_t pos = (uint32_t)(src_i - a.idx_base);
                const uint32_t enc = (a.ins_mode == INS_FIRST) ? (0x7ffffffeu - pos) : (pos + 1u);
                if ((aux & AUX_IDX_MASK) == enc) {
                  const uint32_t sealed = (a.ins_mode == INS_FIRST) ? AUX_SEALED_FIRST : 0u;
                  const uint32_t v = a.vals ? a.vals[src_i] : a.nonexistent;
                  *reinterpret_cast<volatile uint64_t *>(pval) = ((uint64_t)((aux & AUX_STICKY) | sealed) << 32) | v;
                }
              }
            }
            gfound = true;
            act = false;
          } else if ((ge && !gs) || ++probes > bmask) {
            act = false;          // chain ends: absent
          } else {
            b = (b + 1) & bmask;
          }
        }
      }
      // lane gbase + r owns this key's result
      if (gl == r || G == 1) { if (probe) { found = gfound; fval = gval; } }
    }

    if (OP == OP_FIND) { if (valid) a.out_u32[i] = found ? fval : a.nonexistent; }
    if (OP == OP_COUNT) { if (valid) a.out_u8[i] = found ? 1 : 0; }
    if (OP != OP_ERASE) hits += (valid && found) ? 1u : 0u;
  }
  // erase is counted by the lane that won the CAS, everything else by the key's own lane
  hits = warp_sum(hits);
  if (lane == 0 && hits) atomicAdd(a.n_hit, (unsigned long long)hits);
}

// =========================================================================================
// insert
// =========================================================================================
struct InsertArgs {
  TableView t;
  const uint64_t *keys;
  const uint32_t *vals;     // INS_COUNT: per-key addend or null -> dflt.  ordered modes: unused in phase A
  uint64_t n;
  uint32_t dflt;
  int hash_kind;
  HashParams hp;
  uint64_t idx_base;        // ordered modes: batch position of keys[0]
  unsigned long long *n_new;
  int *status;              // set to 1 when a key found no slot (table full)
  // COUNT mode only: a key that finds neither itself nor an empty slot within probe_limit buckets is not
  // placed; its batch index goes to defer_idx (n_new[1] counts them) and the host re-inserts those keys
  // after growing the table.  defer_idx == nullptr: probe the whole table, then report "full".
  uint32_t *defer_idx;
  uint64_t probe_limit;
  // The keys the bulk route's tile insert handed on (kh_bulk.cuh), taken straight from its records and its device-side counter
  // so that no host round trip lies between the two kernels: n_dev != null: the batch has min(n, *n_dev) keys;
  // rec_idx != null: key i = the key of record recs[rec_idx[i] * W] (then `keys` is unused).
  const unsigned long long *n_dev;
  const uint4 *recs;
  const uint32_t *rec_idx;
};

template <int W, int LB, int G, int MODE>
__global__ void __launch_bounds__(256) insert_kernel(InsertArgs a) {
  const uint64_t n_keys = a.n_dev ? min((unsigned long long)a.n, *a.n_dev) : a.n;
  using Ge = Geom<W, LB, G>;
  const int lane = threadIdx.x & 31;
  const int gl = lane & (G - 1);
  const int gbase = lane - gl;
  const uint64_t bmask = a.t.capacity / Ge::BUCKET_SLOTS - 1;
  const int bshift = 64 - log2_u64(bmask + 1);
  const uint64_t warps_total = (uint64_t)gridDim.x * (blockDim.x >> 5);
  const uint64_t warp_id = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  constexpr uint64_t NONE = ~0ULL;
  uint32_t fresh = 0;

  for (uint64_t base = warp_id * 32; base < n_keys; base += warps_total * 32) {
    const uint64_t i = base + lane;
    const bool valid = i < n_keys;
    uint64_t key[W];
#pragma unroll
    for (int q = 0; q < W; ++q) key[q] = 0;
    if (valid) {
      if (a.rec_idx) {
        const uint4 r = __ldg(a.recs + (size_t)a.rec_idx[i] * W);
        key[0] = (uint64_t)r.x | ((uint64_t)r.y << 32);
        if (W == 2) key[W - 1] = (uint64_t)r.z | ((uint64_t)r.w << 32);
      } else load_key<W>(a.keys, i, key);
    }
    uint32_t myv = a.dflt;
    if (MODE == INS_COUNT) { if (valid && a.vals) myv = __ldg(a.vals + i); }
    else {
      const uint32_t pos = (uint32_t)(i - a.idx_base);
      myv = (MODE == INS_FIRST) ? (0x7ffffffeu - pos) : (pos + 1u);   // encoded batch position
    }
    bool probe = valid;
    if (valid && key_is_empty<W>(key)) {     // side slot for the all-ones key
      probe = false;
      if (atomicExch(&a.t.side[0], 1u) == 0u) ++fresh;
      if (MODE == INS_COUNT) atomicAdd(&a.t.side[1], myv); else atomicMax(&a.t.side[2], myv);
    }
    const uint64_t h = top_bits(spread_hash(hash_runtime<W>(a.hash_kind, key, a.hp)), bshift);   // home bucket

#pragma unroll 1
    for (int r = 0; r < G; ++r) {
      uint64_t k[W];
      uint64_t home;
      uint32_t v;
      bool act;
      if (G == 1) {
#pragma unroll
        for (int q = 0; q < W; ++q) k[q] = key[q];
        home = h; act = probe; v = myv;
      } else {
        shfl_key<W>(key, gbase + r, k);
        home = __shfl_sync(FULL_MASK, h, gbase + r);
        act = __shfl_sync(FULL_MASK, (int)probe, gbase + r) != 0;
        v = __shfl_sync(FULL_MASK, myv, gbase + r);
      }
      uint64_t b = home, probes = 0, fe = NONE;
      while (warp_any<G>(act)) {
        LaneScan<W, LB, G> s; s.match = -1; s.empty = -1; s.sticky = false; s.va = 0;
        unsigned char *lane_ptr = a.t.slots + (b * G + gl) * (uint64_t)LB;
        if (act) s = scan_lane<W, LB, G>(lane_ptr, k);
        const uint32_t gm = group_ballot<G>(act && s.match >= 0, gbase);
        const uint32_t ge = group_ballot<G>(act && s.empty >= 0, gbase);
        const uint32_t gs = group_ballot<G>(act && s.sticky, gbase);
        // local index of the first empty slot of the first lane that has one
        int e_local = s.empty;
        if (G > 1) e_local = __shfl_sync(FULL_MASK, s.empty, ge ? (gbase + __ffs(ge) - 1) : lane);
        bool claim = false;
        if (act) {
          if (gm) {
            if (s.match >= 0) {
              uint32_t *pval = reinterpret_cast<uint32_t *>(lane_ptr + s.match * Ge::SLOT + 8 * W);
              if (MODE == INS_COUNT) atomicAdd(pval, v);
              else atomicMax(pval + 1, (((uint32_t)(s.va >> 32)) & AUX_STICKY) | v);
            }
            act = false;
          } else {
            if (fe == NONE && ge) fe = (b * G + (__ffs(ge) - 1)) * Ge::S + e_local;
            if (ge && !gs) claim = true;                      // the chain ends here: the key is absent
            else if (++probes > a.probe_limit) {
              claim = true;
              if (a.defer_idx) fe = NONE;                     // short limit: the key may sit further on -> defer
            } else b = (b + 1) & bmask;
          }
        }
        if (warp_any<G>(claim)) {
          // outcome: 1 claimed, 2 already ours (someone inserted the same key), 3 lost to another key, 4 full
          int outcome = 0;
          if (claim && gl == 0) {
            if (fe == NONE) outcome = 4;
            else {
              unsigned char *slot = a.t.slots + fe * Ge::SLOT;
              uint64_t expect[W], old[W];
#pragma unroll
              for (int q = 0; q < W; ++q) expect[q] = EMPTY_WORD;
              const bool won = cas_key<W>(slot, expect, k, old);
              outcome = won ? 1 : (key_eq<W>(old, k) ? 2 : 3);
              if (outcome != 3) {
                uint32_t *pval = reinterpret_cast<uint32_t *>(slot + 8 * W);
                if (MODE == INS_COUNT) atomicAdd(pval, v);
                else atomicMax(pval + 1, (ldcg_u32(pval + 1) & AUX_STICKY) | v);
              }
            }
          }
          if (G > 1) outcome = __shfl_sync(FULL_MASK, outcome, gbase);
          if (claim) {
            if (outcome == 1) { if (gl == 0) ++fresh; act = false; }
            else if (outcome == 2) act = false;
            else if (outcome == 4) {
              if (gl == 0) {
                if (a.defer_idx) a.defer_idx[atomicAdd(a.n_new + 1, 1ULL)] = (uint32_t)(base + gbase + r);
                else *a.status = 1;
              }
              act = false;
            }
            else { b = home; probes = 0; fe = NONE; }   // lost the race: scan again
          }
        }
      }
    }
  }
  fresh = warp_sum(fresh);
  if (lane == 0 && fresh) atomicAdd(a.n_new, (unsigned long long)fresh);
}

}  // namespace kh
