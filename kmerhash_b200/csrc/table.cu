// table.cu -- kh_table: host side of the HBM k-mer table and its C ABI (include/kmerhash_b200.h).
//
// Host logic mirrors the reference containers' public behaviour:
//   sizing   reserve(n) -> capacity = next_pow2(ceil(n / max_load))      KH/robinhood_offset_hashmap_ptr.hpp:896-981
//   growth   double while the batch would push size past max_load        :1736, :2802-2885 (insert_no_estimate)
//   presize  HLL estimate * (1 + est_error) before the batch             :2512-2535 (insert_impl)
//   shrink   after erase, when size < min_load * capacity                 :3768-3790
// but every data-touching step is a kernel (kh_table.cuh) and a resize is a bulk GPU rebuild.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include <cub/device/device_scan.cuh>
#include <cub/block/block_scan.cuh>

#include "kh_common.h"
#include "kh_table.cuh"
#include "kh_hll.cuh"
#include "kh_partition.h"
#include "kh_window.cuh"
#include "kh_tile.cuh"
#include "kh_table_host.h"

namespace kh {

// ---------------------------------------------------------------------------------------
// small kernels: init, export, seal, compaction helpers
// ---------------------------------------------------------------------------------------
__global__ void init_slots_kernel(uint4 *slots, uint64_t n_units, int words_per_slot_u4) {
  // a slot is 1 (W=1) or 2 (W=2) uint4 units: {key lo, key hi | val, aux, pad}
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_units; i += stride) {
    uint4 v;
    if (words_per_slot_u4 == 1) v = make_uint4(0xffffffffu, 0xffffffffu, 0u, 0u);
    else v = (i & 1) ? make_uint4(0u, 0u, 0u, 0u) : make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
    slots[i] = v;
  }
}

// Compacting dump of the live entries.  A block takes 1024 consecutive slots per step (4 per thread,
// coalesced), ranks the occupied ones with a block scan and reserves its output range with ONE atomicAdd,
// so the output keeps slot order inside a step and steps of the table stay roughly in order (= region
// order, which the rebuild's re-insert relies on for L2 locality).
constexpr int EXPORT_ITEMS = 4;
template <int W>
__global__ void __launch_bounds__(256) export_kernel(TableView t, uint64_t *out_keys, uint32_t *out_vals, uint64_t cap,
                                                     unsigned long long *counter, int include_side) {
  constexpr int SLOT = 16 * W;
  using Scan = cub::BlockScan<uint32_t, 256>;
  __shared__ typename Scan::TempStorage tmp;
  __shared__ unsigned long long block_base;
  const uint64_t step = 256ULL * EXPORT_ITEMS;
  const uint64_t n_steps = (t.capacity + step - 1) / step;
  for (uint64_t st = blockIdx.x; st < n_steps; st += gridDim.x) {
    uint64_t w[EXPORT_ITEMS][2 * W];
    bool occ[EXPORT_ITEMS];
#pragma unroll
    for (int j = 0; j < EXPORT_ITEMS; ++j) {
      const uint64_t i = st * step + (uint64_t)j * 256 + threadIdx.x;
      occ[j] = false;
      if (i < t.capacity) { LdCg<2 * W>::ld(t.slots + i * SLOT, w[j]); occ[j] = !key_is_empty<W>(w[j]); }
    }
    // rank: items are laid out j-major (item j of all threads precedes item j+1), so scan per item row
    uint32_t before[EXPORT_ITEMS], total_rows = 0;
#pragma unroll
    for (int j = 0; j < EXPORT_ITEMS; ++j) {
      uint32_t excl, row_total;
      Scan(tmp).ExclusiveSum(occ[j] ? 1u : 0u, excl, row_total);
      __syncthreads();
      before[j] = total_rows + excl;
      total_rows += row_total;
    }
    if (threadIdx.x == 0) block_base = total_rows ? atomicAdd(counter, (unsigned long long)total_rows) : 0ULL;
    __syncthreads();
    const unsigned long long basepos = block_base;
#pragma unroll
    for (int j = 0; j < EXPORT_ITEMS; ++j) {
      if (occ[j]) {
        const uint64_t pos = basepos + before[j];
        if (pos < cap) {
#pragma unroll
          for (int q = 0; q < W; ++q) out_keys[pos * W + q] = w[j][q];
          if (out_vals) out_vals[pos] = (uint32_t)w[j][W];
        }
      }
    }
    __syncthreads();
  }
  // side slot (all-ones key)
  if (include_side && blockIdx.x == 0 && threadIdx.x == 0 && t.side[0] == 1u) {
    const unsigned long long pos = atomicAdd(counter, 1ULL);
    if (pos < cap) {
      for (int q = 0; q < W; ++q) out_keys[pos * W + q] = EMPTY_WORD;
      if (out_vals) out_vals[pos] = t.side[1];
    }
  }
}

// FIRST tables: every occupied slot carries the "sealed" position so later batches never override it
template <int W>
__global__ void seal_kernel(TableView t) {
  constexpr int SLOT = 16 * W;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < t.capacity; i += stride) {
    uint64_t w[2 * W];
    LdCg<2 * W>::ld(t.slots + i * SLOT, w);
    if (!key_is_empty<W>(w)) {
      uint32_t *paux = reinterpret_cast<uint32_t *>(t.slots + i * SLOT + 8 * W) + 1;
      *paux = (((uint32_t)(w[W] >> 32)) & AUX_STICKY) | AUX_SEALED_FIRST;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0 && t.side[0] == 1u) t.side[2] = AUX_SEALED_FIRST;
}

// stable compaction of the hits of a query batch (find_existing): flags -> positions by scan
template <int W>
__global__ void scatter_hits_kernel(const uint64_t *keys, const uint32_t *vals, const uint8_t *flags,
                                    const uint64_t *pos, uint64_t n, uint64_t *out_keys, uint32_t *out_vals) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (flags[i]) {
      const uint64_t p = pos[i];
#pragma unroll
      for (int q = 0; q < W; ++q) out_keys[p * W + q] = keys[i * W + q];
      out_vals[p] = vals[i];
    }
  }
}
__global__ void widen_flags_kernel(const uint8_t *flags, uint64_t n, uint64_t *out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = flags[i];
}

// AoS -> SoA: std::pair<Kmer, T> records (stride bytes, key at 0, value at val_off, both 8-byte aligned
// record starts) de-interleaved on the device, so the host never touches the batch.
template <int W>
__global__ void split_pairs_kernel(const uint64_t *recs, uint64_t n, uint32_t stride_w, uint32_t val_off_w,
                                   uint32_t val_shift, uint32_t val_mask, uint64_t *keys, uint32_t *vals) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t *r = recs + i * stride_w;
#pragma unroll
    for (int q = 0; q < W; ++q) keys[i * W + q] = r[q];
    vals[i] = (uint32_t)(r[val_off_w] >> val_shift) & val_mask;
  }
}

}  // namespace kh

using namespace kh;

// ---------------------------------------------------------------------------------------
// the table object
// ---------------------------------------------------------------------------------------
namespace {

// KH_TRACE=1: host-side wall time of the stages of an insert / rebuild on stderr (the stream is
// synchronised at every stage boundary while tracing, so the stages do not overlap in the printout)
struct Trace {
  static bool enabled() { static const bool on = getenv("KH_TRACE") && atoi(getenv("KH_TRACE")) != 0; return on; }
  const char *name; cudaStream_t s; std::chrono::steady_clock::time_point t0;
  Trace(const char *name_, cudaStream_t s_) : name(name_), s(s_) {
    if (enabled()) { cudaStreamSynchronize(s); t0 = std::chrono::steady_clock::now(); }
  }
  ~Trace() {
    if (enabled()) {
      cudaStreamSynchronize(s);
      fprintf(stderr, "[kh-trace] %-18s %8.3f ms\n", name, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    }
  }
};

// chunk size (keys) of the host<->device pipeline
constexpr uint64_t HOST_CHUNK = 1ULL << 23;
constexpr uint64_t ORDERED_CHUNK = 1ULL << 30;

#define KH_GEOM_CASE_(Wv, LBv, Gv, ...) \
  { constexpr int W = Wv; constexpr int LB = LBv; constexpr int G = Gv; __VA_ARGS__; }
#define KH_GEOM_SWITCH(t, ...) KH_GEOM_SWITCH3(t, (t)->LB, (t)->G, __VA_ARGS__)
#define KH_GEOM_SWITCH3(t, LBx, Gx, ...)                                                      \
  do {                                                                                        \
    const int _code = (t)->W * 10000 + (LBx) * 100 + (Gx);                                    \
    switch (_code) {                                                                          \
      case 11602: KH_GEOM_CASE_(1, 16, 2, __VA_ARGS__) break;                                 \
      case 11604: KH_GEOM_CASE_(1, 16, 4, __VA_ARGS__) break;                                 \
      case 11608: KH_GEOM_CASE_(1, 16, 8, __VA_ARGS__) break;                                 \
      case 13201: KH_GEOM_CASE_(1, 32, 1, __VA_ARGS__) break;                                 \
      case 13202: KH_GEOM_CASE_(1, 32, 2, __VA_ARGS__) break;                                 \
      case 13204: KH_GEOM_CASE_(1, 32, 4, __VA_ARGS__) break;                                 \
      case 16401: KH_GEOM_CASE_(1, 64, 1, __VA_ARGS__) break;                                 \
      case 26401: KH_GEOM_CASE_(2, 64, 1, __VA_ARGS__) break;                                 \
      case 23201: KH_GEOM_CASE_(2, 32, 1, __VA_ARGS__) break;                                 \
      case 23202: KH_GEOM_CASE_(2, 32, 2, __VA_ARGS__) break;                                 \
      case 23204: KH_GEOM_CASE_(2, 32, 4, __VA_ARGS__) break;                                 \
      default: kh::set_error("unsupported geometry W=%d LB=%d G=%d", (t)->W, (t)->LB, (t)->G); \
               throw kh::CudaError{KH_ERR_INVALID};                                           \
    }                                                                                         \
  } while (0)

bool geometry_ok(int W, int LB, int G) {
  if (LB == 64) return G == 1;
  if (W == 1) return (LB == 16 && (G == 2 || G == 4 || G == 8)) || (LB == 32 && (G == 1 || G == 2 || G == 4));
  return LB == 32 && (G == 1 || G == 2 || G == 4);
}

void default_geometry(int W, int &LB, int &G) {
  LB = 32; G = 2;   // 64-byte buckets, two 32-byte lanes: best insert and find rate in profiles/r01_sweep_probes_geometry.jsonl
  if (const char *e = getenv("KH_GEOM")) {   // "LBxG", experiments only
    int lb = 0, g = 0;
    if (sscanf(e, "%dx%d", &lb, &g) == 2 && geometry_ok(W, lb, g)) { LB = lb; G = g; }
  }
}

void init_slots(kh_table *t, uint64_t cap, unsigned char *p, cudaStream_t s) {
  const uint64_t units = cap * t->slot_bytes() / 16;
  KH_LAUNCH(init_slots_kernel, t->grid_for(units, 256 * 4), 256, 0, s, reinterpret_cast<uint4 *>(p), units, t->W);
}
void alloc_slots(kh_table *t, uint64_t cap, DevBuf<unsigned char> &buf, cudaStream_t s) {
  buf.alloc(cap * t->slot_bytes());
  init_slots(t, cap, buf.p, s);
}
// a table whose initialisation was left to the tile insert (rebuild(..., lazy)) and that takes another path after all
void materialize(kh_table *t, cudaStream_t s) {
  if (!t->uninit) return;
  t->uninit = false;
  init_slots(t, t->capacity, t->slots.p, s);
}

uint64_t capacity_for(const kh_table *t, uint64_t n) {
  uint64_t need = (uint64_t)std::ceil((double)n / t->max_load);
  uint64_t cap = next_pow2(std::max<uint64_t>(need, 1));
  return std::max<uint64_t>(cap, 128);
}

void kev_begin(kh_table *t, int kind, cudaStream_t s) {
  if (!t->kev_on) return;
  if (t->kev_used == t->kev.size()) {
    cudaEvent_t a, b;
    KH_CUDA(cudaEventCreate(&a)); KH_CUDA(cudaEventCreate(&b));
    t->kev.emplace_back(a, b); t->kev_kind.push_back(kind);
  }
  t->kev_kind[t->kev_used] = kind;
  KH_CUDA(cudaEventRecord(t->kev[t->kev_used].first, s));
}
void kev_end(kh_table *t, cudaStream_t s) {
  if (!t->kev_on) return;
  KH_CUDA(cudaEventRecord(t->kev[t->kev_used].second, s));
  ++t->kev_used;
}
void kev_resolve(kh_table *t) {
  for (size_t i = 0; i < t->kev_used; ++i) {
    float ms = 0;
    if (cudaEventSynchronize(t->kev[i].second) == cudaSuccess &&
        cudaEventElapsedTime(&ms, t->kev[i].first, t->kev[i].second) == cudaSuccess) {
      t->kev_ms[t->kev_kind[i]] += ms; t->kev_n[t->kev_kind[i]] += 1;
    }
  }
  t->kev_used = 0;
}

void launch_insert(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n, uint32_t dflt,
                   int mode, cudaStream_t s, uint32_t *defer_idx = nullptr, bool l2_local = false) {
  if (n == 0) return;
  InsertArgs a;
  a.t = t->view(); a.keys = d_keys; a.vals = d_vals; a.n = n; a.dflt = dflt;
  a.hash_kind = t->hash_kind; a.hp = t->hp; a.idx_base = 0;
  a.n_new = t->counters.p; a.status = t->status.p;
  const uint64_t n_buckets = t->capacity / t->bucket_slots();
  a.defer_idx = defer_idx;
  a.probe_limit = defer_idx ? std::min<uint64_t>(n_buckets - 1, 127) : n_buckets - 1;
  const int grid = t->grid_for(n, 256);
  // The table layout is fixed by the BUCKET size (G x LB bytes), not by how a kernel reads a bucket, so a
  // 64-byte-bucket table can be probed 2 lanes x 32 B (best when probes miss the L2: fewer bytes in flight
  // per key, profiles/r01_sweep_*) or 1 thread x 64 B (half the instructions per key: best when the batch
  // is region-sorted and the probes hit the L2).
  int LBk = t->LB, Gk = t->G;
  if (l2_local && t->LB * t->G == 64 && getenv("KH_NO_64x1") == nullptr) { LBk = 64; Gk = 1; }
  kev_begin(t, 0, s);
  KH_GEOM_SWITCH3(t, LBk, Gk, {
    if (mode == INS_COUNT) KH_LAUNCH((insert_kernel<W, LB, G, INS_COUNT>), grid, 256, 0, s, a);
    else if (mode == INS_FIRST) KH_LAUNCH((insert_kernel<W, LB, G, INS_FIRST>), grid, 256, 0, s, a);
    else KH_LAUNCH((insert_kernel<W, LB, G, INS_LAST>), grid, 256, 0, s, a);
  });
  kev_end(t, s);
}

void launch_lookup(kh_table *t, int op, const uint64_t *d_keys, uint64_t n, uint32_t *out_u32, uint8_t *out_u8,
                   const uint32_t *vals, uint32_t nonexistent, unsigned long long *counter, cudaStream_t s) {
  if (n == 0) return;
  LookupArgs a;
  a.t = t->view(); a.keys = d_keys; a.n = n; a.hash_kind = t->hash_kind; a.hp = t->hp;
  a.out_u32 = out_u32; a.out_u8 = out_u8; a.vals = vals; a.nonexistent = nonexistent;
  a.ins_mode = (uint32_t)t->reducer; a.idx_base = 0; a.n_hit = counter;
  const int grid = t->grid_for(n, 256);
  kev_begin(t, 1, s);
  KH_GEOM_SWITCH(t, {
    switch (op) {
      case OP_FIND:     KH_LAUNCH((lookup_kernel<W, LB, G, OP_FIND>), grid, 256, 0, s, a); break;
      case OP_COUNT:    KH_LAUNCH((lookup_kernel<W, LB, G, OP_COUNT>), grid, 256, 0, s, a); break;
      case OP_ERASE:    KH_LAUNCH((lookup_kernel<W, LB, G, OP_ERASE>), grid, 256, 0, s, a); break;
      case OP_UPDATE:   KH_LAUNCH((lookup_kernel<W, LB, G, OP_UPDATE>), grid, 256, 0, s, a); break;
      default:          KH_LAUNCH((lookup_kernel<W, LB, G, OP_FINALIZE>), grid, 256, 0, s, a); break;
    }
  });
  kev_end(t, s);
}

void zero_counter(kh_table *t, cudaStream_t s) {
  KH_CUDA(cudaMemsetAsync(t->counters.p, 0, 2 * sizeof(unsigned long long), s));
  KH_CUDA(cudaMemsetAsync(t->status.p, 0, sizeof(int), s));
}
// reads counter[0] and the status flag; synchronises the stream
uint64_t read_counter(kh_table *t, cudaStream_t s) {
  KH_CUDA(cudaMemcpyAsync(t->h_counters, t->counters.p, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
  KH_CUDA(cudaMemcpyAsync(t->h_status, t->status.p, sizeof(int), cudaMemcpyDeviceToHost, s));
  KH_CUDA(cudaStreamSynchronize(s));
  if (*t->h_status != 0) { set_error("table full: a key found no slot"); throw CudaError{KH_ERR_FULL}; }
  return t->h_counters[0];
}

// compact dump of the live entries into fresh device arrays (used by rebuild and export)
uint64_t dump_entries(kh_table *t, uint64_t *d_keys, uint32_t *d_vals, uint64_t cap, cudaStream_t s, int include_side = 1) {
  KH_CUDA(cudaMemsetAsync(t->counters.p + 1, 0, sizeof(unsigned long long), s));
  TableView v = t->view();
  const int grid = t->grid_for(t->capacity, 256 * 4);
  if (t->W == 1) KH_LAUNCH((export_kernel<1>), grid, 256, 0, s, v, d_keys, d_vals, cap, t->counters.p + 1, include_side);
  else KH_LAUNCH((export_kernel<2>), grid, 256, 0, s, v, d_keys, d_vals, cap, t->counters.p + 1, include_side);
  KH_CUDA(cudaMemcpyAsync(t->h_counters + 1, t->counters.p + 1, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
  KH_CUDA(cudaStreamSynchronize(s));
  return t->h_counters[1];
}

void seal(kh_table *t, cudaStream_t s) {
  TableView v = t->view();
  const int grid = t->grid_for(t->capacity, 256 * 4);
  if (t->W == 1) KH_LAUNCH((seal_kernel<1>), grid, 256, 0, s, v);
  else KH_LAUNCH((seal_kernel<2>), grid, 256, 0, s, v);
}

// bulk rebuild into `new_cap` slots: dump -> fresh table -> insert (keys are unique, fresh vals are 0,
// so the COUNT-mode add is a plain copy for every reducer).  The side slot is not touched.
// lazy: an EMPTY table's new slots are left uninitialised (t->uninit); the caller guarantees that the tile insert, which
// writes every tile of a fresh table anyway, or materialize() follows before anything reads the table.
void rebuild(kh_table *t, uint64_t new_cap, bool lazy = false) {
  cudaStream_t s = t->stream[0];
  Trace tr_all("rebuild", s);
  new_cap = std::max<uint64_t>(next_pow2(new_cap), 128);
  t->has_sticky = false;               // every path below ends with freshly initialised slots
  if (t->capacity == 0 || t->size == 0) {
    t->slots.release();
    t->uninit = false;
    if (lazy) { t->slots.alloc(new_cap * t->slot_bytes()); t->uninit = true; }
    else alloc_slots(t, new_cap, t->slots, s);
    t->capacity = new_cap;
    KH_CUDA(cudaStreamSynchronize(s));
    return;
  }
  materialize(t, s);
  DevBuf<uint64_t> keys(t->size * t->W);
  DevBuf<uint32_t> vals(t->size);
  uint64_t n;
  { Trace tr("rebuild.dump", s); n = dump_entries(t, keys.p, vals.p, t->size, s, 0); }
  { Trace tr("rebuild.alloc+init", s);
    t->slots.release();
    alloc_slots(t, new_cap, t->slots, s);
    t->capacity = new_cap; }
  uint64_t placed;
  { Trace tr("rebuild.insert", s);
    zero_counter(t, s);
    launch_insert(t, keys.p, vals.p, n, 0, INS_COUNT, s, nullptr, true);   // dump order = region order
    placed = read_counter(t, s); }
  if (placed != n) {
    set_error("rebuild lost entries: %llu of %llu", (unsigned long long)placed, (unsigned long long)n);
    throw CudaError{KH_ERR_CUDA};
  }
  if (t->reducer == KH_REDUCE_FIRST) { seal(t, s); KH_CUDA(cudaStreamSynchronize(s)); }
}

void ensure_room(kh_table *t, uint64_t n_more, bool lazy = false) {
  const uint64_t want = capacity_for(t, t->size + n_more);
  if (want > t->capacity) rebuild(t, want, lazy);
}

bool window_candidate(const kh_table *t, const uint32_t *d_vals, uint64_t n) {
  return t->window_on && t->reducer == KH_REDUCE_COUNT && !d_vals && n >= t->window_min_keys && n < (1ULL << 32);
}

// gather the deferred keys (and their addends) of a COUNT launch into fresh arrays
template <int W>
__global__ void gather_deferred_kernel(const uint64_t *keys, const uint32_t *vals, const uint32_t *idx, uint64_t n,
                                       uint64_t *out_keys, uint32_t *out_vals) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint64_t j = idx[i];
#pragma unroll
    for (int q = 0; q < W; ++q) out_keys[i * W + q] = keys[j * W + q];
    if (vals) out_vals[i] = vals[j];
  }
}

// One device-resident piece, any reducer.  Updates t->size and returns #new keys.
// COUNT: a key that finds no slot near its home bucket (table fuller than planned, e.g. a wrong HLL
// estimate) is deferred by the kernel; the table is doubled and the deferred keys are inserted again.
uint64_t insert_resident(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n, uint32_t dflt,
                         cudaStream_t s, bool l2_local = false) {
  uint64_t added = 0;
  const int mode = t->reducer == KH_REDUCE_COUNT ? INS_COUNT : (t->reducer == KH_REDUCE_FIRST ? INS_FIRST : INS_LAST);
  if (mode != INS_COUNT) {
    for (uint64_t off = 0; off < n; off += ORDERED_CHUNK) {
      const uint64_t m = std::min(ORDERED_CHUNK, n - off);
      zero_counter(t, s);
      launch_insert(t, d_keys + off * t->W, d_vals ? d_vals + off : nullptr, m, dflt, mode, s);
      // phase B: the winning batch position stores its value (hits go to counters[1])
      launch_lookup(t, OP_FINALIZE, d_keys + off * t->W, m, nullptr, nullptr, d_vals ? d_vals + off : nullptr, dflt,
                    t->counters.p + 1, s);
      const uint64_t got = read_counter(t, s);
      added += got; t->size += got;
    }
    return added;
  }
  constexpr uint64_t COUNT_CHUNK = 1ULL << 31;     // deferred indices are 32-bit
  for (uint64_t off = 0; off < n; off += COUNT_CHUNK) {
    uint64_t m = std::min(COUNT_CHUNK, n - off);
    const uint64_t *ck = d_keys + off * t->W;
    const uint32_t *cv = d_vals ? d_vals + off : nullptr;
    DevBuf<uint64_t> rk;          // deferred keys / addends of the previous round
    DevBuf<uint32_t> rv;
    while (m) {
      DevBuf<uint32_t> defer(m);
      zero_counter(t, s);
      launch_insert(t, ck, cv, m, dflt, INS_COUNT, s, defer.p, l2_local);
      const uint64_t got = read_counter(t, s);
      const uint64_t n_def = t->h_counters[1];
      added += got; t->size += got;
      if (n_def == 0) break;
      DevBuf<uint64_t> nk(n_def * t->W);
      DevBuf<uint32_t> nv(cv ? n_def : 0);
      if (t->W == 1) KH_LAUNCH((gather_deferred_kernel<1>), t->grid_for(n_def, 256), 256, 0, s, ck, cv, defer.p, n_def, nk.p, nv.p);
      else KH_LAUNCH((gather_deferred_kernel<2>), t->grid_for(n_def, 256), 256, 0, s, ck, cv, defer.p, n_def, nk.p, nv.p);
      KH_CUDA(cudaStreamSynchronize(s));
      rk = std::move(nk); rv = std::move(nv);
      ck = rk.p; cv = cv ? rv.p : nullptr; m = n_def;
      rebuild(t, t->capacity * 2);
    }
  }
  return added;
}

// L2-windowed bulk insert.  Random probes into a table much larger than the 126 MB L2 run at the DRAM
// random-access rate (~13 G read+atomic/s measured, profiles/r01_probes_l2_resident.jsonl); into a region
// that stays resident in L2 they run at ~55 G/s.  So a large COUNT batch is first sorted by table region
// (kh_window.cuh: top bits of the spread store hash = a contiguous 1/P of the table, window_bytes each),
// and the probe kernel then walks the sorted batch: at any moment all resident warps work on one or two
// regions.  Costs one extra streaming pass (16 B/key), saves the DRAM-latency-bound random traffic.
// `d_fine`: the fine region histogram of this batch if the HLL pass already produced it, else nullptr.
// Shared-memory tile insert (kh_tile.cuh): eligible for COUNT batches with a uniform addend into an 8-byte-key
// table of 64-byte buckets without sticky bits, when the batch is large against the table (the whole table is
// streamed through shared memory once) and a coarse region spans at most 8 tiles.
bool tile_candidate(const kh_table *t, const uint32_t *d_vals, uint64_t n, int &lp, int &subs_log2) {
  static const bool on = !(getenv("KH_TILE") && atoi(getenv("KH_TILE")) == 0);
  if (!on || !window_candidate(t, d_vals, n)) return false;
  if (t->W != 1 || t->LB * t->G != 64 || t->has_sticky || t->capacity < (uint64_t)TI_SLOTS) return false;
  if (n * 4 < t->capacity) return false;
  const int lbk = log2_u64(t->capacity / 4);               // bits of a bucket index
  lp = std::min(REGION_FINE_BITS, lbk - TI_BUCKET_BITS);
  subs_log2 = lbk - TI_BUCKET_BITS - lp;
  return subs_log2 <= 3;
}

// a batch handed over in pieces (SegSpec) -> one contiguous array, for the paths whose kernels want one
const uint64_t *flatten_segments(kh_table *t, const uint64_t *d_keys, uint64_t n, const SegSpec *segs, DevBuf<uint64_t> &buf, cudaStream_t s) {
  if (!segs || segs->nseg <= 1) return d_keys;
  buf.alloc(n * t->W);
  for (int q = 0; q < segs->nseg; ++q) {
    const uint64_t c = segs->prefix[q + 1] - segs->prefix[q];
    if (c) KH_CUDA(cudaMemcpyAsync(buf.p + segs->prefix[q] * t->W, d_keys + segs->base[q] * t->W, c * 8 * t->W, cudaMemcpyDeviceToDevice, s));
  }
  return buf.p;
}

uint64_t insert_tiled(kh_table *t, const uint64_t *d_keys, uint64_t n, uint32_t dflt, const unsigned long long *d_fine,
                      int lp, int subs_log2, cudaStream_t s, const SegSpec *segs = nullptr) {
  KH_CUDA(cudaFuncSetAttribute(tile_insert_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TI_SMEM));   // per device
  static const bool old_scatter = getenv("KH_SCATTER") && !strcmp(getenv("KH_SCATTER"), "old");
  constexpr int CSTRIDE = 16;          // one 128-byte line per cursor
  DevBuf<unsigned long long> fine_own, cursors((size_t)CSTRIDE << lp);
  DevBuf<uint64_t> flat;
  if (!d_fine || old_scatter) { d_keys = flatten_segments(t, d_keys, n, segs, flat, s); segs = nullptr; }
  if (!d_fine) {
    fine_own.alloc(REGION_FINE);
    kev_begin(t, 3, s);
    hll_estimate_dev(t->hash_kind, t->hp, t->W, d_keys, n, nullptr, s, t->sms, fine_own.p);
    kev_end(t, s);
    d_fine = fine_own.p;
  }
  KH_LAUNCH(window_cursor_kernel, 1, REGION_FINE, 0, s, d_fine, lp, cursors.p, CSTRIDE);
  DevBuf<uint64_t> sorted(n);
  DevBuf<uint16_t> tags(((n + TI_CHUNK - 1) / TI_CHUNK + 1) * TI_CHUNK);
  const int lbk = log2_u64(t->capacity / 4);
  { Trace tr("tile.scatter", s);
    const BinSpec spec{0, lp, 0u, t->hash_kind, t->hp};
    const TagSpec ts{tags.p, 64 - lbk, (1u << (lbk - lp)) - 1u, CSTRIDE};
    kev_begin(t, 2, s);
    if (old_scatter) KH_LAUNCH((window_scatter_kernel<1, false, false, true>), t->sms * 4, WIN_BLOCK, 0, s, d_keys, n, spec, cursors.p,
                               sorted.p, nullptr, BinDst{}, ts);
    else region_scatter_launch<1, true>(t->sms, s, d_keys, n, spec, cursors.p, sorted.p, ts, segs);
    kev_end(t, s); }
  DevBuf<uint32_t> defer(n);
  uint64_t added = 0, n_def = 0;
  { Trace tr("tile.insert", s);
    zero_counter(t, s);
    TileArgs a;
    a.t = t->view(); a.keys = sorted.p; a.tags = tags.p; a.bin_end = cursors.p; a.bin_stride = CSTRIDE; a.subs_log2 = subs_log2;
    a.fresh = t->uninit ? 2 : (t->size == 0 ? 1 : 0); a.dflt = dflt;   // 2: the slots hold garbage, every tile is written
    a.tags32 = nullptr; a.slab = 0; a.lbk = lbk;
    t->uninit = false; a.n_new = t->counters.p; a.defer_idx = defer.p;
    const uint64_t n_tiles = t->capacity / TI_SLOTS;
    const int grid = (int)std::min<uint64_t>(n_tiles, (uint64_t)t->sms * 2);
    kev_begin(t, 0, s);
    KH_LAUNCH(tile_insert_kernel<false>, grid, TI_THREADS, TI_SMEM, s, a);
    kev_end(t, s);
    added = read_counter(t, s);
    n_def = t->h_counters[1];
    t->size += added; }
  if (n_def) {       // chains that leave their tile, the all-ones key: a handful of keys through the global kernel
    Trace tr("tile.deferred", s);
    DevBuf<uint64_t> nk(n_def);
    KH_LAUNCH((gather_deferred_kernel<1>), t->grid_for(n_def, 256), 256, 0, s, sorted.p, (const uint32_t *)nullptr, defer.p, n_def,
              nk.p, (uint32_t *)nullptr);
    added += insert_resident(t, nk.p, nullptr, n_def, dflt, s);
  }
  return added;
}

uint64_t insert_windowed(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n, uint32_t dflt,
                         const unsigned long long *d_fine, cudaStream_t s, const SegSpec *segs = nullptr);

// One-read bulk insert of a batch whose distinct count is not known yet (the reference's insert(): estimate -> reserve ->
// insert, KH/robinhood_offset_hashmap_ptr.hpp:2512-2535).  The region scatter runs FIRST, in SLAB mode: it needs no
// histogram (every one of the 1024 regions owns a fixed range sized for a uniform hash, the rare overflow goes to a side
// list), so the same pass that sorts the batch also computes the HyperLogLog registers.  Then: estimate -> reserve (slots
// left uninitialised) -> tile insert cutting its keys out of the 32-bit hash tags -> deferred and overflow keys through the
// global kernel.  Returns false without touching the table when the batch does not qualify (the caller takes the
// two-pass route); when only the table turns out unsuitable after the estimate, the region counts serve as the histogram
// of the two-pass route, so nothing is computed twice but the sort.
bool fused_candidate(const kh_table *t, const uint32_t *d_vals, uint64_t n) {
  static const bool on = !(getenv("KH_FUSED") && atoi(getenv("KH_FUSED")) == 0);
  static const bool tile_on = !(getenv("KH_TILE") && atoi(getenv("KH_TILE")) == 0);
  if (!on || !tile_on || !window_candidate(t, d_vals, n)) return false;
  // n < 2^30: output positions are 32-bit in the kernel; even a batch that falls into ONE region keeps its cursor
  // (1023 * slab + n < 2.2 n) below 2^32
  if (t->W != 1 || t->LB * t->G != 64 || t->has_sticky || n >= (1ULL << 30) || n < (1ULL << 20)) return false;
  return t->capacity <= (1ULL << 26);          // a larger table has more than 8 tiles per region: no tile insert anyway
}

// state of one fused batch: begin -> feed (once, or chunk by chunk as the keys arrive over PCIe) -> finish
struct FusedBatch {
  static constexpr int CSTRIDE = 16;
  uint64_t n = 0, slab = 0, ovf_cap = 0;
  DevBuf<unsigned long long> cursors, ovf_count;
  DevBuf<uint64_t> sorted, ovf;
  DevBuf<uint32_t> tags32, regs;
};
void fused_begin(FusedBatch &f, kh_table *t, uint64_t n, cudaStream_t s) {
  f.n = n;
  f.slab = ((uint64_t)((double)n / REGION_FINE * 1.125) + 2048 + 127) & ~127ULL;
  f.ovf_cap = n / 16 + 4096;
  f.cursors.alloc((size_t)FusedBatch::CSTRIDE * REGION_FINE); f.ovf_count.alloc(1);
  f.sorted.alloc(f.slab * REGION_FINE); f.ovf.alloc(f.ovf_cap);
  f.tags32.alloc(f.slab * REGION_FINE + 2 * TI_CHUNK); f.regs.alloc(HLL_M);
  KH_CUDA(cudaMemsetAsync(f.regs.p, 0, HLL_M * sizeof(uint32_t), s));
  KH_CUDA(cudaMemsetAsync(f.ovf_count.p, 0, sizeof(unsigned long long), s));
  KH_LAUNCH(slab_cursor_kernel, (WIN_MAX_P + 255) / 256, 256, 0, s, f.cursors.p, FusedBatch::CSTRIDE, (unsigned long long)f.slab);
}
void fused_feed(FusedBatch &f, kh_table *t, const uint64_t *d_keys, uint64_t m, cudaStream_t s) {
  if (!m) return;
  const BinSpec spec{0, REGION_FINE_BITS, 0u, t->hash_kind, t->hp};
  SlabSpec sl{};
  sl.slab = f.slab; sl.ovf_keys = f.ovf.p; sl.ovf_count = f.ovf_count.p; sl.ovf_cap = f.ovf_cap; sl.g_regs = f.regs.p; sl.tags32 = f.tags32.p;
  sl.h32 = hash_bits(t->hash_kind) == 32 ? 1 : 0;
  kev_begin(t, 2, s);
  region_scatter_slab_launch(t->sms, s, d_keys, m, spec, f.cursors.p, FusedBatch::CSTRIDE, f.sorted.p, sl);
  kev_end(t, s);
}

// d_keys: the whole batch (only read again when the two-pass route has to take over)
int64_t fused_finish(FusedBatch &f, kh_table *t, const uint64_t *d_keys, uint32_t dflt, cudaStream_t s) {
  constexpr int CSTRIDE = FusedBatch::CSTRIDE;
  const double err = 1.0 + hll_error_rate();
  const uint64_t size0 = t->size, n = f.n, slab = f.slab, ovf_cap = f.ovf_cap;
  DevBuf<unsigned long long> &cursors = f.cursors, &ovf_count = f.ovf_count;
  DevBuf<uint64_t> &sorted = f.sorted, &ovf = f.ovf;
  DevBuf<uint32_t> &tags32 = f.tags32, &regs = f.regs;
  unsigned long long n_ovf = 0;
  KH_CUDA(cudaMemcpyAsync(&n_ovf, ovf_count.p, sizeof(n_ovf), cudaMemcpyDeviceToHost, s));
  const double est = hll_finish(t->hash_kind, regs.p, nullptr, s);       // synchronises s
  const bool lazy = t->size == 0;
  ensure_room(t, (uint64_t)std::ceil(std::min((double)n, est * err)), lazy);
  int lp = 0, subs = 0;
  if (n_ovf > ovf_cap || !tile_candidate(t, nullptr, n, lp, subs)) {
    // the table does not suit the tile insert (or the batch is heavily skewed): two-pass route, histogram = region counts
    std::vector<unsigned long long> h((size_t)CSTRIDE * REGION_FINE), fine_h(REGION_FINE);
    KH_CUDA(cudaMemcpyAsync(h.data(), cursors.p, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    KH_CUDA(cudaStreamSynchronize(s));
    for (int r = 0; r < REGION_FINE; ++r) fine_h[r] = h[(size_t)r * CSTRIDE] - (unsigned long long)r * slab;
    DevBuf<unsigned long long> fine(REGION_FINE);
    KH_CUDA(cudaMemcpyAsync(fine.p, fine_h.data(), REGION_FINE * sizeof(unsigned long long), cudaMemcpyHostToDevice, s));
    sorted.release(); tags32.release(); ovf.release();
    insert_windowed(t, d_keys, nullptr, n, dflt, fine.p, s);
    return (int64_t)(t->size - size0);
  }
  KH_CUDA(cudaFuncSetAttribute(tile_insert_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TI_SMEM));   // per device
  DevBuf<uint32_t> defer(n);
  uint64_t n_def = 0;
  { Trace tr("tile.insert", s);
    zero_counter(t, s);
    TileArgs a;
    a.t = t->view(); a.keys = sorted.p; a.tags = nullptr; a.bin_end = cursors.p; a.bin_stride = CSTRIDE; a.subs_log2 = 0;
    a.fresh = t->uninit ? 2 : (t->size == 0 ? 1 : 0); a.dflt = dflt; a.n_new = t->counters.p; a.defer_idx = defer.p;
    a.tags32 = tags32.p; a.slab = slab; a.lbk = log2_u64(t->capacity / 4);
    t->uninit = false;
    const uint64_t n_tiles = t->capacity / TI_SLOTS;
    const int grid = (int)std::min<uint64_t>(n_tiles, (uint64_t)t->sms * 2);
    kev_begin(t, 0, s);
    KH_LAUNCH(tile_insert_kernel<true>, grid, TI_THREADS, TI_SMEM, s, a);
    kev_end(t, s);
    t->size += read_counter(t, s);
    n_def = t->h_counters[1]; }
  if (n_def) {
    Trace tr("tile.deferred", s);
    DevBuf<uint64_t> nk(n_def);
    KH_LAUNCH((gather_deferred_kernel<1>), t->grid_for(n_def, 256), 256, 0, s, sorted.p, (const uint32_t *)nullptr, defer.p, n_def,
              nk.p, (uint32_t *)nullptr);
    insert_resident(t, nk.p, nullptr, n_def, dflt, s);
  }
  if (n_ovf) insert_resident(t, ovf.p, nullptr, n_ovf, dflt, s);
  KH_CUDA(cudaStreamSynchronize(s));
  return (int64_t)(t->size - size0);
}

int64_t insert_fused(kh_table *t, const uint64_t *d_keys, uint64_t n, uint32_t dflt, cudaStream_t s) {
  FusedBatch f;
  fused_begin(f, t, n, s);
  { Trace tr("fused.hll+scatter", s); fused_feed(f, t, d_keys, n, s); }
  return fused_finish(f, t, d_keys, dflt, s);
}

uint64_t insert_windowed(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n, uint32_t dflt,
                         const unsigned long long *d_fine, cudaStream_t s, const SegSpec *segs) {
  int tile_lp = 0, tile_subs = 0;
  if (tile_candidate(t, d_vals, n, tile_lp, tile_subs)) return insert_tiled(t, d_keys, n, dflt, d_fine, tile_lp, tile_subs, s, segs);
  materialize(t, s);
  const uint64_t table_bytes = t->capacity * (uint64_t)t->slot_bytes();
  const bool windowed = window_candidate(t, d_vals, n) && table_bytes >= 4 * t->window_bytes;
  DevBuf<uint64_t> flat;
  if (!windowed || !d_fine) { d_keys = flatten_segments(t, d_keys, n, segs, flat, s); segs = nullptr; }
  if (!windowed) return insert_resident(t, d_keys, d_vals, n, dflt, s);
  const int lp = std::min(log2_u64(next_pow2(table_bytes / t->window_bytes)), REGION_FINE_BITS);
  DevBuf<unsigned long long> fine_own, cursors(1u << lp);
  if (!d_fine) {
    fine_own.alloc(REGION_FINE);
    hll_estimate_dev(t->hash_kind, t->hp, t->W, d_keys, n, nullptr, s, t->sms, fine_own.p);
    d_fine = fine_own.p;
  }
  KH_LAUNCH(window_cursor_kernel, 1, REGION_FINE, 0, s, d_fine, lp, cursors.p);
  DevBuf<uint64_t> sorted(n * t->W);
  { Trace tr("window.scatter", s);
    const BinSpec spec{0, lp, 0u, t->hash_kind, t->hp};
    const TagSpec ts{nullptr, 32, 0u, 1};
    if (t->W == 1) region_scatter_launch<1, false>(t->sms, s, d_keys, n, spec, cursors.p, sorted.p, ts, segs);
    else region_scatter_launch<2, false>(t->sms, s, d_keys, n, spec, cursors.p, sorted.p, ts, segs); }
  Trace tr("window.insert", s);
  return insert_resident(t, sorted.p, nullptr, n, dflt, s, true);   // synchronises s before the scratch is released
}

}  // namespace

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
extern "C" {

kh_table *kh_create(int device, int map_kind, int key_words, int hash_kind, uint32_t seed, int reducer) {
  kh_table *t = nullptr;
  int64_t rc = guarded([&]() -> int64_t {
    if (key_words < 1 || key_words > 2) { set_error("key_words must be 1 or 2"); return KH_ERR_INVALID; }
    if (!hash_kind_supported(hash_kind)) { set_error("hash kind %d is not defined for 8/16-byte keys", hash_kind); return KH_ERR_INVALID; }
    if (reducer < 0 || reducer > 2 || map_kind < 0 || map_kind > 2) { set_error("bad reducer / map kind"); return KH_ERR_INVALID; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    if (device < 0 || device >= ndev) { set_error("device %d out of range", device); return KH_ERR_INVALID; }
    DeviceGuard g(device);
    t = new kh_table();
    t->device = device; t->map_kind = map_kind; t->W = key_words;
    t->hash_kind = fold_kind(hash_kind); t->seed = seed; t->reducer = reducer;
    switch (map_kind) {   // reference defaults
      case KH_MAP_ROBINHOOD:   t->min_load = 0.4; t->max_load = 0.9; break;
      case KH_MAP_LINEARPROBE: t->min_load = 0.2; t->max_load = 0.6; break;
      default:                 t->min_load = 0.2; t->max_load = 0.8; break;
    }
    default_geometry(t->W, t->LB, t->G);
    if (const char *e = getenv("KH_WINDOW")) t->window_on = atoi(e) != 0;
    t->hp = make_hash_params(t->hash_kind, seed, key_words);
    t->sms = sm_count(device);
    for (int i = 0; i < 2; ++i) {
      KH_CUDA(cudaStreamCreateWithFlags(&t->stream[i], cudaStreamNonBlocking));
      KH_CUDA(cudaEventCreateWithFlags(&t->ev[i], cudaEventDisableTiming));
    }
    t->side.alloc(4); t->counters.alloc(2); t->status.alloc(1);
    KH_CUDA(cudaMemset(t->side.p, 0, 4 * sizeof(uint32_t)));
    KH_CUDA(cudaHostAlloc((void **)&t->h_counters, 2 * sizeof(unsigned long long), cudaHostAllocDefault));
    KH_CUDA(cudaHostAlloc((void **)&t->h_status, sizeof(int), cudaHostAllocDefault));
    rebuild(t, 128);
    return KH_OK;
  });
  if (rc != KH_OK) { if (t) { kh_destroy(t); } return nullptr; }
  return t;
}

void kh_destroy(kh_table *t) {
  if (!t) return;
  int prev = -1; cudaGetDevice(&prev); cudaSetDevice(t->device);
  for (int i = 0; i < 2; ++i) { if (t->stream[i]) cudaStreamDestroy(t->stream[i]); if (t->ev[i]) cudaEventDestroy(t->ev[i]); }
  for (auto &e : t->kev) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
  if (t->h_counters) cudaFreeHost(t->h_counters);
  if (t->h_status) cudaFreeHost(t->h_status);
  t->slots.release(); t->side.release(); t->counters.release(); t->status.release();
  for (int i = 0; i < 2; ++i) { t->d_keys[i].release(); t->d_vals[i].release(); t->d_u8[i].release(); }
  delete t;
  if (prev >= 0) cudaSetDevice(prev);
}

int kh_set_load_factors(kh_table *t, double min_load, double max_load) {
  if (!t || !(max_load > 0.0 && max_load < 1.0) || !(min_load >= 0.0 && min_load < max_load)) {
    set_error("load factors must satisfy 0 <= min < max < 1"); return KH_ERR_INVALID;
  }
  t->min_load = min_load; t->max_load = max_load;
  return KH_OK;
}
int kh_set_lookahead(kh_table *t, int, int) { return t ? KH_OK : KH_ERR_INVALID; }

void *kh_stream(kh_table *t) { return t ? (void *)t->stream[0] : nullptr; }

int kh_kernel_timing(kh_table *t, int enable) {
  if (!t) return KH_ERR_INVALID;
  t->kev_on = enable != 0; t->kev_used = 0;
  for (int i = 0; i < 4; ++i) { t->kev_ms[i] = 0.0; t->kev_n[i] = 0; }
  return KH_OK;
}
int kh_kernel_time(kh_table *t, int which, double *sum_ms, uint64_t *launches) {
  if (!t || which < 0 || which > 3) return KH_ERR_INVALID;
  int prev = -1; cudaGetDevice(&prev); cudaSetDevice(t->device);
  kev_resolve(t);
  if (prev >= 0) cudaSetDevice(prev);
  if (sum_ms) *sum_ms = t->kev_ms[which];
  if (launches) *launches = t->kev_n[which];
  return KH_OK;
}

int kh_set_insert_window(kh_table *t, int enable, uint64_t window_bytes, uint64_t min_keys) {
  if (!t || (enable && (window_bytes < 1024 || (window_bytes & (window_bytes - 1))))) {
    set_error("window_bytes must be a power of two >= 1024"); return KH_ERR_INVALID;
  }
  t->window_on = enable != 0;
  if (enable) { t->window_bytes = window_bytes; t->window_min_keys = min_keys; }
  return KH_OK;
}

int kh_set_geometry(kh_table *t, int lane_bytes, int lanes_per_bucket) {
  return (int)guarded([&]() -> int64_t {
    if (!t || !geometry_ok(t->W, lane_bytes, lanes_per_bucket)) { set_error("unsupported geometry"); return KH_ERR_INVALID; }
    if (t->size != 0) { set_error("geometry can only change on an empty table"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    t->LB = lane_bytes; t->G = lanes_per_bucket;
    return KH_OK;
  });
}

int kh_reserve(kh_table *t, uint64_t n) {
  return (int)guarded([&]() -> int64_t {
    if (!t) return KH_ERR_INVALID;
    DeviceGuard g(t->device);
    const uint64_t want = capacity_for(t, n);
    if (want > t->capacity) rebuild(t, want);
    return KH_OK;
  });
}
int kh_rehash(kh_table *t, uint64_t slots) {
  return (int)guarded([&]() -> int64_t {
    if (!t) return KH_ERR_INVALID;
    DeviceGuard g(t->device);
    uint64_t want = std::max(next_pow2(slots), capacity_for(t, t->size));
    if (want != t->capacity) rebuild(t, want);
    return KH_OK;
  });
}
uint64_t kh_capacity(const kh_table *t) { return t ? t->capacity : 0; }
uint64_t kh_size(const kh_table *t) { return t ? t->size : 0; }
double kh_load_factor(const kh_table *t) { return (t && t->capacity) ? (double)t->size / (double)t->capacity : 0.0; }
int kh_device(const kh_table *t) { return t ? t->device : -1; }
int kh_key_words(const kh_table *t) { return t ? t->W : 0; }

int kh_clear(kh_table *t) {
  return (int)guarded([&]() -> int64_t {
    if (!t) return KH_ERR_INVALID;
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    const uint64_t units = t->capacity * t->slot_bytes() / 16;
    KH_LAUNCH(init_slots_kernel, t->grid_for(units, 256 * 4), 256, 0, s, reinterpret_cast<uint4 *>(t->slots.p), units, t->W);
    KH_CUDA(cudaMemsetAsync(t->side.p, 0, 4 * sizeof(uint32_t), s));
    KH_CUDA(cudaStreamSynchronize(s));
    t->size = 0;
    t->has_sticky = false;
    return KH_OK;
  });
}

// ---- HLL ---------------------------------------------------------------------------------
double kh_estimate_distinct_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, uint8_t *regs_out) {
  double est = -1.0;
  guarded([&]() -> int64_t {
    if (!t) return KH_ERR_INVALID;
    DeviceGuard g(t->device);
    est = hll_estimate_dev(t->hash_kind, t->hp, t->W, d_keys, n, regs_out, t->stream[0], t->sms);
    return KH_OK;
  });
  return est;
}

double kh_estimate_distinct_host(kh_table *t, const uint64_t *keys, uint64_t n, uint8_t *regs_out) {
  double est = -1.0;
  guarded([&]() -> int64_t {
    if (!t) return KH_ERR_INVALID;
    DeviceGuard g(t->device);
    DevBuf<uint64_t> d(std::max<uint64_t>(n * t->W, 1));
    KH_CUDA(cudaMemcpy(d.p, keys, n * t->W * sizeof(uint64_t), cudaMemcpyHostToDevice));
    est = hll_estimate_dev(t->hash_kind, t->hp, t->W, d.p, n, regs_out, t->stream[0], t->sms);
    return KH_OK;
  });
  return est;
}

}  // extern "C"

// ---- insert ------------------------------------------------------------------------------
// Growth policy without an estimate = the reference's insert_no_estimate: fill up to max_load,
// double, continue (KH/robinhood_offset_hashmap_ptr.hpp:2802-2885).
int64_t kh::insert_dev_impl(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n,
                            uint32_t dflt, int estimate, const PreEstimate *pre) {
  cudaStream_t s = t->stream[0];
  if (n == 0) return 0;
  const double err = 1.0 + hll_error_rate();
  const uint64_t size0 = t->size;
  // the HLL pass also histograms the batch by table region when the windowed insert may follow
  DevBuf<unsigned long long> fine;
  bool have_fine = false;
  auto estimate_batch = [&]() -> double {
    if (pre) return pre->est;
    Trace tr("hll+hist", s);
    if (window_candidate(t, d_vals, n)) { fine.alloc(REGION_FINE); have_fine = true; }
    kev_begin(t, 3, s);
    const double e = hll_estimate_dev(t->hash_kind, t->hp, t->W, d_keys, n, nullptr, s, t->sms, have_fine ? fine.p : nullptr);
    kev_end(t, s);
    return e;
  };
  // Can the whole batch go in as ONE piece?  Yes when even its distinct keys (an HLL bound) keep the load legal.
  // (COUNT tolerates a wrong bound: keys that find no slot are deferred and re-inserted after growing.)
  bool single = false;
  const bool count_mode = t->reducer == KH_REDUCE_COUNT;
  if (estimate && !pre && fused_candidate(t, d_vals, n)) {
    Trace tr_ins("insert.total(fused)", s);
    struct Materialize {
      kh_table *t; cudaStream_t s;
      ~Materialize() { if (t->uninit) { try { materialize(t, s); cudaStreamSynchronize(s); } catch (...) {} } }
    } guard{t, s};
    insert_fused(t, d_keys, n, dflt, s);
    while ((double)t->size > t->max_load * (double)t->capacity) rebuild(t, t->capacity * 2);
    return (int64_t)(t->size - size0);
  }
  if (estimate) {
    // insert_impl: reserve(size + estimate * (1 + error)) then one pass (:2512-2535)
    const double est = estimate_batch();
    // an empty COUNT table about to take a large uniform batch: the tile insert writes every tile, skip the init pass
    const bool lazy = t->size == 0 && window_candidate(t, d_vals, n);
    ensure_room(t, (uint64_t)std::ceil(std::min((double)n, est * err)), lazy);
    single = count_mode;
  }
  struct Materialize {                  // whatever path follows (or throws): never leave uninitialised slots behind
    kh_table *t; cudaStream_t s;
    ~Materialize() { if (t->uninit) { try { materialize(t, s); cudaStreamSynchronize(s); } catch (...) {} } }
  } mat_guard{t, s};
  if (!single) materialize(t, s);
  if (!single && (double)(t->size + n) <= t->max_load * (double)t->capacity) single = true;
  if (!single && !estimate && count_mode && n >= (1ULL << 16)) {
    // insert_no_estimate on a table the caller reserved (the drivers: estimate -> reserve -> insert_no_estimate):
    // the HLL pass costs one streaming read and saves the piecewise schedule below
    const double est = estimate_batch();
    single = (double)t->size + std::min((double)n, est * err) <= t->max_load * (double)t->capacity;
  }
  Trace tr_ins("insert.total", s);
  const SegSpec *segs = pre ? pre->segs : nullptr;
  DevBuf<uint64_t> flat;
  if (single) {
    insert_windowed(t, d_keys, d_vals, n, dflt, pre ? pre->fine : (have_fine ? fine.p : nullptr), s, segs);
  } else {
    d_keys = flatten_segments(t, d_keys, n, segs, flat, s);
    // Growth without an estimate = the reference's insert_no_estimate: fill up to max_load, double, continue.
    // A piece may overshoot max_load (duplicates usually keep it far below); it never fills the table:
    // the hard limit is halfway between max_load and 1.
    const double hard = 0.5 * (t->max_load + 1.0);
    uint64_t off = 0;
    while (off < n) {
      uint64_t room = (uint64_t)std::floor(hard * (double)t->capacity);
      room = room > t->size ? room - t->size : 0;
      if (room == 0 || (double)t->size >= t->max_load * (double)t->capacity) { rebuild(t, t->capacity * 2); continue; }
      const uint64_t m = std::min(n - off, room);
      insert_resident(t, d_keys + off * t->W, d_vals ? d_vals + off : nullptr, m, dflt, s);
      off += m;
    }
  }
  while ((double)t->size > t->max_load * (double)t->capacity) rebuild(t, t->capacity * 2);
  return (int64_t)(t->size - size0);
}

extern "C" {

int64_t kh_insert_dev(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n,
                      uint32_t default_val, int estimate) {
  return guarded([&]() -> int64_t {
    if (!t || (!d_keys && n)) { set_error("null table or keys"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    return insert_dev_impl(t, d_keys, d_vals, n, default_val, estimate);
  });
}

static bool is_pinned(const void *p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost;
}

int64_t kh_insert_host(kh_table *t, const uint64_t *keys, const uint32_t *vals, uint64_t n,
                       uint32_t default_val, int estimate) {
  return guarded([&]() -> int64_t {
    if (!t || (!keys && n)) { set_error("null table or keys"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    if (n == 0) return 0;
    const int W = t->W;
    if (estimate && !vals && fused_candidate(t, nullptr, n)) {
      // One-read route, fed chunk by chunk BEHIND the PCIe copy: when the last chunk lands, the batch is already sorted by
      // table region and its HLL registers are complete; only the estimate, the reserve and the tile insert remain.
      DevBuf<uint64_t> dk(n * W);
      FusedBatch f;
      fused_begin(f, t, n, t->stream[0]);
      int b = 0;
      for (uint64_t off = 0; off < n; off += HOST_CHUNK, b ^= 1) {
        const uint64_t m = std::min(HOST_CHUNK, n - off);
        KH_CUDA(cudaMemcpyAsync(dk.p + off * W, keys + off * W, m * W * 8, cudaMemcpyHostToDevice, t->stream[1]));
        KH_CUDA(cudaEventRecord(t->ev[b], t->stream[1]));
        KH_CUDA(cudaStreamWaitEvent(t->stream[0], t->ev[b], 0));
        fused_feed(f, t, dk.p + off * W, m, t->stream[0]);
      }
      KH_CUDA(cudaStreamSynchronize(t->stream[1]));
      const uint64_t size0 = t->size;
      struct Materialize {
        kh_table *t; cudaStream_t s;
        ~Materialize() { if (t->uninit) { try { materialize(t, s); cudaStreamSynchronize(s); } catch (...) {} } }
      } guard{t, t->stream[0]};
      fused_finish(f, t, dk.p, default_val, t->stream[0]);
      while ((double)t->size > t->max_load * (double)t->capacity) rebuild(t, t->capacity * 2);
      return (int64_t)(t->size - size0);
    }
    if (estimate || t->reducer != KH_REDUCE_COUNT) {
      // Whole batch resident (the ordered reducers need stable batch positions, and presizing needs the
      // estimate before the first insert).  The HLL registers and the region histogram are accumulated
      // chunk by chunk BEHIND the PCIe copy, so the estimate is ready when the last chunk lands.
      DevBuf<uint64_t> dk(n * W);
      DevBuf<uint32_t> dv(vals ? n : 0);
      const bool pipe_hll = estimate != 0;
      DevBuf<uint32_t> regs(pipe_hll ? HLL_M : 0);
      DevBuf<unsigned long long> fine(pipe_hll && window_candidate(t, vals ? dv.p : nullptr, n) ? REGION_FINE : 0);
      if (pipe_hll) hll_reset(regs.p, fine.p, t->stream[0]);
      int b = 0;
      for (uint64_t off = 0; off < n; off += HOST_CHUNK, b ^= 1) {
        const uint64_t m = std::min(HOST_CHUNK, n - off);
        KH_CUDA(cudaMemcpyAsync(dk.p + off * W, keys + off * W, m * W * 8, cudaMemcpyHostToDevice, t->stream[1]));
        if (vals) KH_CUDA(cudaMemcpyAsync(dv.p + off, vals + off, m * 4, cudaMemcpyHostToDevice, t->stream[1]));
        if (pipe_hll) {
          KH_CUDA(cudaEventRecord(t->ev[b], t->stream[1]));
          KH_CUDA(cudaStreamWaitEvent(t->stream[0], t->ev[b], 0));
          hll_accumulate(t->hash_kind, t->hp, W, dk.p + off * W, m, regs.p, fine.p, t->stream[0], t->sms);
        }
      }
      KH_CUDA(cudaStreamSynchronize(t->stream[1]));
      if (!pipe_hll) return insert_dev_impl(t, dk.p, vals ? dv.p : nullptr, n, default_val, 0);
      PreEstimate pre{hll_finish(t->hash_kind, regs.p, nullptr, t->stream[0]), fine.p};
      return insert_dev_impl(t, dk.p, vals ? dv.p : nullptr, n, default_val, estimate, &pre);
    }
    // COUNT without estimate: stream chunks; chunk i+1 is copied while chunk i is inserted
    uint64_t added = 0;
    const bool pinned = is_pinned(keys) && (!vals || is_pinned(vals));
    (void)pinned;   // pageable sources still work (the driver stages them); pinned ones overlap
    for (int b = 0; b < 2; ++b) { t->d_keys[b].ensure(HOST_CHUNK * W); if (vals) t->d_vals[b].ensure(HOST_CHUNK); }
    uint64_t off = 0; int buf = 0;
    uint64_t pending_m[2] = {0, 0};
    auto issue_copy = [&](int b, uint64_t o, uint64_t m) {
      KH_CUDA(cudaMemcpyAsync(t->d_keys[b].p, keys + o * W, m * W * 8, cudaMemcpyHostToDevice, t->stream[1]));
      if (vals) KH_CUDA(cudaMemcpyAsync(t->d_vals[b].p, vals + o, m * 4, cudaMemcpyHostToDevice, t->stream[1]));
      KH_CUDA(cudaEventRecord(t->ev[b], t->stream[1]));
      pending_m[b] = m;
    };
    issue_copy(0, 0, std::min(HOST_CHUNK, n));
    while (off < n) {
      const uint64_t m = pending_m[buf];
      const uint64_t next_off = off + m;
      // the other buffer is free: its insert was synchronised in the previous iteration
      if (next_off < n) issue_copy(buf ^ 1, next_off, std::min(HOST_CHUNK, n - next_off));
      KH_CUDA(cudaStreamWaitEvent(t->stream[0], t->ev[buf], 0));
      const int64_t got = insert_dev_impl(t, t->d_keys[buf].p, vals ? t->d_vals[buf].p : nullptr, m, default_val, 0);
      if (got < 0) return got;
      added += (uint64_t)got;
      off = next_off; buf ^= 1;
    }
    return (int64_t)added;
  });
}

int64_t kh_insert_pairs_host(kh_table *t, const void *pairs, uint64_t n, uint32_t stride_bytes,
                             uint32_t val_offset, uint32_t val_bytes, int estimate) {
  return guarded([&]() -> int64_t {
    if (!t || (!pairs && n)) { set_error("null table or pairs"); return KH_ERR_INVALID; }
    const int W = t->W;
    if (stride_bytes % 8 || stride_bytes < (uint32_t)(8 * W) + val_bytes || val_offset < (uint32_t)(8 * W) ||
        val_offset + val_bytes > stride_bytes || !(val_bytes == 1 || val_bytes == 2 || val_bytes == 4) ||
        (val_offset % 8) + val_bytes > 8) {
      set_error("unsupported pair layout: stride %u, value %u bytes at %u", stride_bytes, val_bytes, val_offset);
      return KH_ERR_INVALID;
    }
    DeviceGuard g(t->device);
    if (n == 0) return 0;
    const uint32_t stride_w = stride_bytes / 8, val_off_w = val_offset / 8, val_shift = 8 * (val_offset % 8);
    const uint32_t val_mask = val_bytes == 4 ? 0xffffffffu : ((1u << (8 * val_bytes)) - 1u);
    DevBuf<uint64_t> dk(n * W);
    DevBuf<uint32_t> dv(n);
    DevBuf<uint64_t> stage[2];
    const uint64_t chunk = HOST_CHUNK / 2;
    for (int b = 0; b < 2; ++b) stage[b].alloc(std::min(chunk, n) * stride_w);
    int buf = 0;
    for (uint64_t off = 0; off < n; off += chunk, buf ^= 1) {
      const uint64_t m = std::min(chunk, n - off);
      cudaStream_t s = t->stream[buf];   // copy-in and split of one staging buffer are ordered on its stream
      KH_CUDA(cudaMemcpyAsync(stage[buf].p, (const unsigned char *)pairs + off * stride_bytes, m * stride_bytes,
                              cudaMemcpyHostToDevice, s));
      if (W == 1) KH_LAUNCH((split_pairs_kernel<1>), t->grid_for(m, 1024), 256, 0, s, stage[buf].p, m, stride_w, val_off_w, val_shift, val_mask, dk.p + off, dv.p + off);
      else KH_LAUNCH((split_pairs_kernel<2>), t->grid_for(m, 1024), 256, 0, s, stage[buf].p, m, stride_w, val_off_w, val_shift, val_mask, dk.p + off * 2, dv.p + off);
    }
    KH_CUDA(cudaStreamSynchronize(t->stream[1]));
    KH_CUDA(cudaStreamSynchronize(t->stream[0]));
    return insert_dev_impl(t, dk.p, dv.p, n, 0, estimate);
  });
}

}  // extern "C"

// ---- queries -----------------------------------------------------------------------------
template <typename OutT, typename Launch>
static int64_t query_host(kh_table *t, const uint64_t *keys, uint64_t n, OutT *out, DevBuf<OutT> *dout, Launch launch) {
  // double-buffered: H2D(keys i+1) | kernel(i) | D2H(results i-1)
  const int W = t->W;
  if (n == 0) return 0;
  for (int b = 0; b < 2; ++b) { t->d_keys[b].ensure(HOST_CHUNK * W); dout[b].ensure(HOST_CHUNK); }
  zero_counter(t, t->stream[0]);
  KH_CUDA(cudaStreamSynchronize(t->stream[0]));
  cudaStream_t s2[2] = {t->stream[0], t->stream[1]};
  int buf = 0;
  for (uint64_t off = 0; off < n; off += HOST_CHUNK, buf ^= 1) {
    const uint64_t m = std::min(HOST_CHUNK, n - off);
    cudaStream_t s = s2[buf];   // each buffer has its own stream: copy-in, kernel, copy-out in order
    KH_CUDA(cudaMemcpyAsync(t->d_keys[buf].p, keys + off * W, m * W * 8, cudaMemcpyHostToDevice, s));
    launch(t->d_keys[buf].p, m, dout[buf].p, s);
    KH_CUDA(cudaMemcpyAsync(out + off, dout[buf].p, m * sizeof(OutT), cudaMemcpyDeviceToHost, s));
  }
  KH_CUDA(cudaStreamSynchronize(t->stream[1]));
  return (int64_t)read_counter(t, t->stream[0]);
}

extern "C" {

int64_t kh_count_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, uint8_t *d_out) {
  return guarded([&]() -> int64_t {
    if (!t || (n && (!d_keys || !d_out))) { set_error("null argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    zero_counter(t, s);
    launch_lookup(t, OP_COUNT, d_keys, n, nullptr, d_out, nullptr, 0, t->counters.p, s);
    return (int64_t)read_counter(t, s);
  });
}
int64_t kh_count_host(kh_table *t, const uint64_t *keys, uint64_t n, uint8_t *out) {
  return guarded([&]() -> int64_t {
    if (!t || (n && (!keys || !out))) { set_error("null argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    return query_host<uint8_t>(t, keys, n, out, t->d_u8, [&](const uint64_t *dk, uint64_t m, uint8_t *dout, cudaStream_t s) {
      launch_lookup(t, OP_COUNT, dk, m, nullptr, dout, nullptr, 0, t->counters.p, s);
    });
  });
}
int64_t kh_find_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, uint32_t *d_out, uint32_t nonexistent) {
  return guarded([&]() -> int64_t {
    if (!t || (n && (!d_keys || !d_out))) { set_error("null argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    zero_counter(t, s);
    launch_lookup(t, OP_FIND, d_keys, n, d_out, nullptr, nullptr, nonexistent, t->counters.p, s);
    return (int64_t)read_counter(t, s);
  });
}
int64_t kh_find_host(kh_table *t, const uint64_t *keys, uint64_t n, uint32_t *out, uint32_t nonexistent) {
  return guarded([&]() -> int64_t {
    if (!t || (n && (!keys || !out))) { set_error("null argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    return query_host<uint32_t>(t, keys, n, out, t->d_vals, [&](const uint64_t *dk, uint64_t m, uint32_t *dout, cudaStream_t s) {
      launch_lookup(t, OP_FIND, dk, m, dout, nullptr, nullptr, nonexistent, t->counters.p, s);
    });
  });
}

int64_t kh_find_existing_dev(kh_table *t, const uint64_t *d_keys, uint64_t n, uint64_t *d_out_keys, uint32_t *d_out_vals) {
  return guarded([&]() -> int64_t {
    if (!t || (n && (!d_keys || !d_out_keys || !d_out_vals))) { set_error("null argument"); return KH_ERR_INVALID; }
    if (n == 0) return 0;
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    DevBuf<uint32_t> vals(n);
    DevBuf<uint8_t> flags(n);
    DevBuf<uint64_t> pos(n), wide(n);
    zero_counter(t, s);
    launch_lookup(t, OP_FIND, d_keys, n, vals.p, nullptr, nullptr, 0, t->counters.p, s);
    launch_lookup(t, OP_COUNT, d_keys, n, nullptr, flags.p, nullptr, 0, t->counters.p, s);
    KH_LAUNCH(widen_flags_kernel, t->grid_for(n, 1024), 256, 0, s, flags.p, n, wide.p);
    size_t tmp_bytes = 0;
    KH_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, wide.p, pos.p, n, s));
    DevBuf<unsigned char> tmp(tmp_bytes);
    KH_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, wide.p, pos.p, n, s));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (t->W == 1) KH_LAUNCH((scatter_hits_kernel<1>), t->grid_for(n, 1024), 256, 0, s, d_keys, vals.p, flags.p, pos.p, n, d_out_keys, d_out_vals);
    else KH_LAUNCH((scatter_hits_kernel<2>), t->grid_for(n, 1024), 256, 0, s, d_keys, vals.p, flags.p, pos.p, n, d_out_keys, d_out_vals);
    return (int64_t)(read_counter(t, s) / 2);   // both lookups counted the hits
  });
}
int64_t kh_find_existing_host(kh_table *t, const uint64_t *keys, uint64_t n, uint64_t *out_keys, uint32_t *out_vals) {
  return guarded([&]() -> int64_t {
    if (!t || (n && (!keys || !out_keys || !out_vals))) { set_error("null argument"); return KH_ERR_INVALID; }
    if (n == 0) return 0;
    DeviceGuard g(t->device);
    DevBuf<uint64_t> dk(n * t->W), ok(n * t->W);
    DevBuf<uint32_t> ov(n);
    KH_CUDA(cudaMemcpy(dk.p, keys, n * t->W * 8, cudaMemcpyHostToDevice));
    const int64_t hits = kh_find_existing_dev(t, dk.p, n, ok.p, ov.p);
    if (hits > 0) {
      KH_CUDA(cudaMemcpy(out_keys, ok.p, (uint64_t)hits * t->W * 8, cudaMemcpyDeviceToHost));
      KH_CUDA(cudaMemcpy(out_vals, ov.p, (uint64_t)hits * 4, cudaMemcpyDeviceToHost));
    }
    return hits;
  });
}

int64_t kh_update_dev(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n) {
  return guarded([&]() -> int64_t {
    if (!t || (n && (!d_keys || !d_vals))) { set_error("null argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    zero_counter(t, s);
    launch_lookup(t, OP_UPDATE, d_keys, n, nullptr, nullptr, d_vals, 0, t->counters.p, s);
    return (int64_t)read_counter(t, s);
  });
}
int64_t kh_update_host(kh_table *t, const uint64_t *keys, const uint32_t *vals, uint64_t n) {
  return guarded([&]() -> int64_t {
    if (!t || (n && (!keys || !vals))) { set_error("null argument"); return KH_ERR_INVALID; }
    if (n == 0) return 0;
    DeviceGuard g(t->device);
    DevBuf<uint64_t> dk(n * t->W); DevBuf<uint32_t> dv(n);
    KH_CUDA(cudaMemcpy(dk.p, keys, n * t->W * 8, cudaMemcpyHostToDevice));
    KH_CUDA(cudaMemcpy(dv.p, vals, n * 4, cudaMemcpyHostToDevice));
    return kh_update_dev(t, dk.p, dv.p, n);
  });
}

// ---- erase -------------------------------------------------------------------------------
static void maybe_shrink(kh_table *t) {
  // the reference reserves for the surviving size once the load drops below min_load (:3787)
  if ((double)t->size < t->min_load * (double)t->capacity) {
    const uint64_t want = capacity_for(t, t->size);
    if (want < t->capacity) rebuild(t, want);
  }
}
int64_t kh_erase_dev(kh_table *t, const uint64_t *d_keys, uint64_t n) {
  return guarded([&]() -> int64_t {
    if (!t || (n && !d_keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    cudaStream_t s = t->stream[0];
    uint64_t erased;
    { Trace tr("erase.kernel", s);
      zero_counter(t, s);
      launch_lookup(t, OP_ERASE, d_keys, n, nullptr, nullptr, nullptr, 0, t->counters.p, s);
      erased = read_counter(t, s); }
    t->size -= erased;
    if (erased) t->has_sticky = true;
    maybe_shrink(t);
    return (int64_t)erased;
  });
}
int64_t kh_erase_host(kh_table *t, const uint64_t *keys, uint64_t n) {
  return guarded([&]() -> int64_t {
    if (!t || (n && !keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    if (n == 0) return 0;
    DeviceGuard g(t->device);
    const int W = t->W;
    for (int b = 0; b < 2; ++b) t->d_keys[b].ensure(HOST_CHUNK * W);
    zero_counter(t, t->stream[0]);
    KH_CUDA(cudaStreamSynchronize(t->stream[0]));
    int buf = 0;
    for (uint64_t off = 0; off < n; off += HOST_CHUNK, buf ^= 1) {
      const uint64_t m = std::min(HOST_CHUNK, n - off);
      cudaStream_t s = t->stream[buf];
      KH_CUDA(cudaMemcpyAsync(t->d_keys[buf].p, keys + off * W, m * W * 8, cudaMemcpyHostToDevice, s));
      launch_lookup(t, OP_ERASE, t->d_keys[buf].p, m, nullptr, nullptr, nullptr, 0, t->counters.p, s);
    }
    KH_CUDA(cudaStreamSynchronize(t->stream[1]));
    const uint64_t erased = read_counter(t, t->stream[0]);
    t->size -= erased;
    if (erased) t->has_sticky = true;
    maybe_shrink(t);
    return (int64_t)erased;
  });
}

// ---- export ------------------------------------------------------------------------------
int64_t kh_export_dev(kh_table *t, uint64_t *d_out_keys, uint32_t *d_out_vals, uint64_t cap) {
  return guarded([&]() -> int64_t {
    if (!t || (cap && !d_out_keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    return (int64_t)dump_entries(t, d_out_keys, d_out_vals, cap, t->stream[0]);
  });
}
int64_t kh_export_host(kh_table *t, uint64_t *out_keys, uint32_t *out_vals, uint64_t cap) {
  return guarded([&]() -> int64_t {
    if (!t || (cap && !out_keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    const uint64_t m = std::min<uint64_t>(cap, t->size);
    DevBuf<uint64_t> dk(std::max<uint64_t>(m * t->W, 1)); DevBuf<uint32_t> dv(std::max<uint64_t>(m, 1));
    const uint64_t n = dump_entries(t, dk.p, dv.p, m, t->stream[0]);
    const uint64_t c = std::min(n, m);
    if (c) {
      KH_CUDA(cudaMemcpy(out_keys, dk.p, c * t->W * 8, cudaMemcpyDeviceToHost));
      if (out_vals) KH_CUDA(cudaMemcpy(out_vals, dv.p, c * 4, cudaMemcpyDeviceToHost));
    }
    return (int64_t)n;
  });
}

}  // extern "C"
