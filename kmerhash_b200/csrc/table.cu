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
#include <cub/device/device_radix_sort.cuh>
#include <cub/block/block_scan.cuh>

#include "kh_common.h"
#include "kh_table.cuh"
#include "kh_hll.cuh"
#include "kh_partition.h"
#include "kh_bulk.cuh"
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
  t->kev_kind[t->kev_used] = t->kev_in_rebuild ? 4 : kind;
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

struct RecSource { const unsigned long long *n_dev; const uint4 *recs; const uint32_t *rec_idx; };   // InsertArgs::n_dev ...
void launch_insert(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n, uint32_t dflt,
                   int mode, cudaStream_t s, uint32_t *defer_idx = nullptr, bool l2_local = false, const RecSource *src = nullptr) {
  if (n == 0) return;
  InsertArgs a;
  a.n_dev = src ? src->n_dev : nullptr; a.recs = src ? src->recs : nullptr; a.rec_idx = src ? src->rec_idx : nullptr;
  a.t = t->view(); a.keys = d_keys; a.vals = d_vals; a.n = n; a.dflt = dflt;
  a.hash_kind = t->hash_kind; a.hp = t->hp; a.idx_base = 0;
  a.n_new = t->counters.p; a.status = t->status.p;
  const uint64_t n_buckets = t->capacity / t->bucket_slots();
  a.defer_idx = defer_idx;
  a.probe_limit = defer_idx ? std::min<uint64_t>(n_buckets - 1, 127) : n_buckets - 1;
  const int grid = src ? t->sms : t->grid_for(n, 256);              // src: n is an upper bound, the true count is small
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
  KH_CUDA(cudaMemsetAsync(t->counters.p, 0, 3 * sizeof(unsigned long long), s));   // counters and status
}
// reads counter[0] and the status flag; synchronises the stream
uint64_t read_counter(kh_table *t, cudaStream_t s) {
  KH_CUDA(cudaMemcpyAsync(t->h_counters, t->counters.p, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
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

// Streaming rebuild (kh_bulk.cuh: tile_rebuild_kernel): new tile x is fed from the contiguous range of old slots that hash
// into it, the old table is read once and the new one written once.  Entries that sit outside "their" range (a chain that
// crossed a tile boundary) or whose chain leaves the new tile come back as a short (key, value) list for the global kernel.
// Returns false -- the old table untouched -- when the geometry does not fit or that list overflowed.
bool rebuild_streamed(kh_table *t, uint64_t new_cap, cudaStream_t s) {
  static const bool on = !(getenv("KH_BULK") && atoi(getenv("KH_BULK")) == 0);
  const uint64_t tile_slots = (uint64_t)BK_TILE_BYTES / t->slot_bytes();
  if (!on || t->LB * t->G != 64 || new_cap < tile_slots || t->capacity < tile_slots) return false;
  const int W = t->W;
  DevBuf<unsigned char> fresh(new_cap * t->slot_bytes());          // every tile is written by the kernel
  const uint64_t dcap = t->size / 8 + 4096;
  DevBuf<uint64_t> dk(dcap * W);
  DevBuf<uint32_t> dv(dcap);
  uint64_t placed, n_def;
  uint32_t side_present = 0;             // the all-ones key lives in the side slot, not in the slots
  { Trace tr("rebuild.stream", s);
    zero_counter(t, s);
    RebuildArgs a;
    a.src = t->view(); a.dst = TableView{fresh.p, new_cap, t->side.p};
    a.hash_kind = t->hash_kind; a.hp = t->hp; a.lbk = log2_u64(new_cap / t->bucket_slots());
    a.n_placed = t->counters.p; a.defer_keys = dk.p; a.defer_vals = dv.p; a.defer_cap = dcap;
    const int grid = (int)std::min<uint64_t>(new_cap / tile_slots, (uint64_t)t->sms * 2);
    kev_begin(t, 0, s);
    if (W == 1) {
      KH_CUDA(cudaFuncSetAttribute(tile_rebuild_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rebuild_smem_bytes<1>()));
      KH_LAUNCH(tile_rebuild_kernel<1>, grid, TI_THREADS, rebuild_smem_bytes<1>(), s, a);
    } else {
      KH_CUDA(cudaFuncSetAttribute(tile_rebuild_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rebuild_smem_bytes<2>()));
      KH_LAUNCH(tile_rebuild_kernel<2>, grid, TI_THREADS, rebuild_smem_bytes<2>(), s, a);
    }
    kev_end(t, s);
    KH_CUDA(cudaMemcpyAsync(&side_present, t->side.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    placed = read_counter(t, s);
    n_def = t->h_counters[1]; }
  if (n_def > dcap) return false;
  t->slots = std::move(fresh);
  t->capacity = new_cap;
  if (n_def) {
    Trace tr("rebuild.deferred", s);
    zero_counter(t, s);
    launch_insert(t, dk.p, dv.p, n_def, 0, INS_COUNT, s);
    placed += read_counter(t, s);
  }
  if (placed + (side_present == 1u ? 1u : 0u) != t->size) {
    set_error("rebuild lost entries: %llu of %llu", (unsigned long long)placed, (unsigned long long)t->size);
    throw CudaError{KH_ERR_CUDA};
  }
  return true;
}

// bulk rebuild into `new_cap` slots.  Large tables: one streaming pass (rebuild_streamed).  Small ones: dump -> fresh table ->
// insert (keys are unique, fresh vals are 0, so the COUNT-mode add is a plain copy for every reducer).  The side slot is not touched.
// lazy: an EMPTY table's new slots are left uninitialised (t->uninit); the caller guarantees that the tile insert, which
// writes every tile of a fresh table anyway, or materialize() follows before anything reads the table.
void rebuild(kh_table *t, uint64_t new_cap, bool lazy = false) {
  cudaStream_t s = t->stream[0];
  Trace tr_all("rebuild", s);
  struct InRebuild { kh_table *t; bool prev; InRebuild(kh_table *t_) : t(t_), prev(t_->kev_in_rebuild) { t->kev_in_rebuild = true; } ~InRebuild() { t->kev_in_rebuild = prev; } } in_rebuild(t);
  new_cap = std::max<uint64_t>(next_pow2(new_cap), 128);
  if (t->capacity == 0 || t->size == 0) {
    t->slots.release();
    t->uninit = false;
    if (lazy) { t->slots.alloc(new_cap * t->slot_bytes()); t->uninit = true; }
    else alloc_slots(t, new_cap, t->slots, s);
    t->capacity = new_cap;
    t->has_sticky = false;
    KH_CUDA(cudaStreamSynchronize(s));
    return;
  }
  materialize(t, s);
  if (!rebuild_streamed(t, new_cap, s)) {
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
  }
  t->has_sticky = false;               // freshly written slots carry no sticky bit
  if (t->reducer == KH_REDUCE_FIRST) { seal(t, s); KH_CUDA(cudaStreamSynchronize(s)); }
}

// ---- presizing: the reference keeps ONE HyperLogLog over everything ever inserted with an estimate and reserves for ITS
// estimate (KH/robinhood_offset_hashmap_ptr.hpp:291,2512-2535); reserve -> rehash resizes -- up or down -- whenever the target
// differs and the entries fit (:896-901, :907-935).
double merge_batch_registers(kh_table *t, const uint8_t *batch_regs) {
  for (int i = 0; i < HLL_M; ++i) t->hll[i] = std::max(t->hll[i], batch_regs[i]);
  return hll_estimate_from_regs(t->hll, hash_bits(t->hash_kind));
}
void reserve_for_estimate(kh_table *t, double est, bool lazy) {
  const uint64_t want = capacity_for(t, (uint64_t)(est * (1.0 + hll_error_rate())));
  if (want != t->capacity && (double)t->size < std::ceil(t->max_load * (double)want)) rebuild(t, want, lazy);
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

// =========================================================================================================================
// the streaming bulk insert (kh_bulk.cuh)
// =========================================================================================================================
bool bulk_on() { static const bool on = !(getenv("KH_BULK") && atoi(getenv("KH_BULK")) == 0); return on; }
constexpr uint64_t BULK_MAX_KEYS = 3ULL << 29;      // record positions are 32-bit: (1.125 + 1) * n + pads stays below 2^32

// COUNT batch with one addend for all keys, or a FIRST / LAST batch (values or not) that lies in one piece; 64-byte buckets,
// large enough to pay for streaming the table once
bool bulk_candidate(const kh_table *t, const uint32_t *d_vals, uint64_t n, const BkSegs *segs = nullptr) {
  static const bool ord_on = !(getenv("KH_BULK_ORD") && atoi(getenv("KH_BULK_ORD")) == 0);
  const bool kind_ok = t->reducer == KH_REDUCE_COUNT ? !d_vals : (ord_on && !segs);
  return bulk_on() && t->window_on && kind_ok && t->LB * t->G == 64 && n >= t->window_min_keys && n >= 1024 && n <= BULK_MAX_KEYS;
}

// level-1 regions: as many as the LARGEST table this batch can lead to has tiles (then region == tile or finer and no
// level 2 is needed), at most 4096, at least 64
int bulk_region_bits(const kh_table *t, uint64_t n) {
  const uint64_t tile_slots = (uint64_t)BK_TILE_BYTES / t->slot_bytes();
  const uint64_t worst = std::max(t->capacity, capacity_for(t, t->size + n));
  int lp = log2_u64(std::max<uint64_t>(worst / tile_slots, 1));
  if (const char *e = getenv("KH_BULK_LP")) lp = atoi(e);          // experiments / tests: force level 2
  return std::max(6, std::min(lp, bk_l1_max_lp(t->W)));
}

void bulk_begin(BulkBatch &b, kh_table *t, uint64_t n, cudaStream_t s) {
  b.n = n; b.fed = 0;
  b.lp = bulk_region_bits(t, n);
  const uint64_t R = 1ULL << b.lp;
  // a region's share of a uniformly hashed batch plus 12.5 % and a pad of 8 standard deviations of small shares
  b.slab = ((uint64_t)((double)n / (double)R * 1.125) + 1024 + 127) & ~127ULL;
  b.ovf_cap = n / 16 + 4096;
  { size_t off = 0;
    auto take = [&](size_t bytes) { const size_t at = off; off += (bytes + 255) & ~(size_t)255; return at; };
    const size_t o_recs = take(b.slab * R * t->W * sizeof(uint4)), o_ovf = take(b.ovf_cap * t->W * 8), o_cur = take((size_t)BK_CSTRIDE * R * 8),
                 o_regs = take(HLL_M * 4 + 8), o_cnt = o_regs + HLL_M * 4, o_def = take(n * 4);   // registers | overflow count: one memset, one read-back
    b.arena.alloc(off);
    b.recs.p = reinterpret_cast<uint4 *>(b.arena.p + o_recs); b.ovf.p = reinterpret_cast<uint64_t *>(b.arena.p + o_ovf);
    b.cursors.p = reinterpret_cast<unsigned long long *>(b.arena.p + o_cur); b.ovf_count.p = reinterpret_cast<unsigned long long *>(b.arena.p + o_cnt);
    b.regs.p = reinterpret_cast<uint32_t *>(b.arena.p + o_regs); b.defer.p = reinterpret_cast<uint32_t *>(b.arena.p + o_def); }
  KH_CUDA(cudaMemsetAsync(b.regs.p, 0, HLL_M * sizeof(uint32_t) + sizeof(unsigned long long), s));
  KH_LAUNCH(l1_cursor_kernel, (int)((R + 255) / 256), 256, 0, s, b.cursors.p, (int)R, (unsigned long long)b.slab);
}

// m keys at d_keys (or in the pieces of `segs`) -> level 1.  Asynchronous on s.
// dsegs != nullptr: the pieces lie in a DEVICE-resident table (written by a kernel earlier on s); m is then only an upper
// bound, the grid is fixed and b.fed is set by the caller once the counts are known
void bulk_feed(BulkBatch &b, kh_table *t, const uint64_t *d_keys, uint64_t m, cudaStream_t s, const BkSegs *segs = nullptr,
               const BkSegs *dsegs = nullptr) {
  if (!m) return;
  L1Args a;
  a.pos_base = 0;
  a.keys = d_keys; a.n = m; a.lp = b.lp; a.hash_kind = t->hash_kind; a.h32 = hash_bits(t->hash_kind) == 32 ? 1 : 0; a.hp = t->hp;
  a.cursors = b.cursors.p; a.recs = b.recs.p; a.slab = b.slab;
  a.ovf_keys = b.ovf.p; a.ovf_count = b.ovf_count.p; a.ovf_cap = b.ovf_cap; a.g_regs = b.regs.p;
  const uint64_t tile = l1_tile_keys(t->W);
  BkSegs sg{};
  sg.nseg = 1;
  uint64_t n_tiles = (m + tile - 1) / tile;
  if (segs && segs->nseg > 1) {
    sg = *segs;
    sg.tprefix[0] = 0;
    for (int q = 0; q < sg.nseg; ++q) sg.tprefix[q + 1] = sg.tprefix[q] + (sg.prefix[q + 1] - sg.prefix[q] + tile - 1) / tile;
    n_tiles = sg.tprefix[sg.nseg];
  } else if (segs && segs->nseg == 1) {
    a.keys = d_keys + segs->base[0] * t->W;
  }
  if (!n_tiles) return;
  // dsegs (the pieces of a pipelined exchange): one SM stays free for the one-warp barrier kernel of the next piece, which could
  // not start while these persistent CTAs -- each holds a whole SM's registers -- are running
  const int grid = dsegs ? std::max(t->sms - 1, 1) : (int)std::min<uint64_t>(n_tiles, (uint64_t)t->sms);
  kev_begin(t, 2, s);
  if (b.ord) {                                                      // records with batch positions (one piece)
    a.pos_base = b.fed;
    if (t->W == 1) {
      KH_CUDA(cudaFuncSetAttribute(l1_scatter_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L1Cfg<1>::SMEM_ORD));
      KH_LAUNCH((l1_scatter_kernel<1, true>), grid, L1_THREADS, L1Cfg<1>::SMEM_ORD, s, a, sg, dsegs);
    } else {
      KH_CUDA(cudaFuncSetAttribute(l1_scatter_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L1Cfg<2>::SMEM_ORD));
      KH_LAUNCH((l1_scatter_kernel<2, true>), grid, L1_THREADS, L1Cfg<2>::SMEM_ORD, s, a, sg, dsegs);
    }
  } else if (t->W == 1) {
    KH_CUDA(cudaFuncSetAttribute(l1_scatter_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L1Cfg<1>::SMEM));   // per device
    KH_LAUNCH(l1_scatter_kernel<1>, grid, L1_THREADS, L1Cfg<1>::SMEM, s, a, sg, dsegs);
  } else {
    KH_CUDA(cudaFuncSetAttribute(l1_scatter_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L1Cfg<2>::SMEM));
    KH_LAUNCH(l1_scatter_kernel<2>, grid, L1_THREADS, L1Cfg<2>::SMEM, s, a, sg, dsegs);
  }
  kev_end(t, s);
  if (!dsegs) b.fed += m;
}

// a batch handed over in pieces -> one contiguous array, for the paths whose kernels want one
const uint64_t *flatten_segments(kh_table *t, const uint64_t *d_keys, uint64_t n, const BkSegs *segs, DevBuf<uint64_t> &buf, cudaStream_t s) {
  if (!segs || segs->nseg < 1) return d_keys;
  if (segs->nseg == 1) return d_keys + segs->base[0] * t->W;
  buf.alloc(n * t->W);
  for (int q = 0; q < segs->nseg; ++q) {
    const uint64_t c = segs->prefix[q + 1] - segs->prefix[q];
    if (c) KH_CUDA(cudaMemcpyAsync(buf.p + segs->prefix[q] * t->W, d_keys + segs->base[q] * t->W, c * 8 * t->W, cudaMemcpyDeviceToDevice, s));
  }
  return buf.p;
}

// the general route for a resident batch: fill up to max_load, double, continue (the reference's insert_no_estimate,
// KH/robinhood_offset_hashmap_ptr.hpp:2802-2885); `single`: the caller knows the batch fits
void insert_general(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n, uint32_t dflt, bool single, cudaStream_t s) {
  materialize(t, s);
  if (!single && (double)(t->size + n) <= t->max_load * (double)t->capacity) single = true;
  if (single) { insert_resident(t, d_keys, d_vals, n, dflt, s); return; }
  // A piece may overshoot max_load (duplicates usually keep it far below); it never fills the table: the hard limit is
  // halfway between max_load and 1.
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

// estimate -> reserve -> [level 2] -> tile insert -> deferred and overflow keys.  d_keys / segs: the whole batch again, read
// only when the general route has to take over (skewed batch, table that does not suit the tiles).
// estimate != 0: insert(): the batch registers join the table's HyperLogLog and the table is reserved for its estimate.
// estimate == 0: insert_no_estimate(): the table keeps its capacity when the batch fits, else the general route grows it.
int64_t bulk_finish(BulkBatch &b, kh_table *t, const uint64_t *d_keys, const BkSegs *segs, uint32_t dflt, int estimate, cudaStream_t s) {
  const uint64_t size0 = t->size, n = b.n;
  const int W = t->W;
  // registers and overflow count: one copy into pinned memory (a pageable target would be staged by the driver)
  KH_CUDA(cudaMemcpyAsync(t->h_regs, b.regs.p, HLL_M * sizeof(uint32_t) + sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
  KH_CUDA(cudaStreamSynchronize(s));
  unsigned long long n_ovf = 0;
  std::memcpy(&n_ovf, t->h_regs + HLL_M, sizeof(n_ovf));
  uint8_t batch_regs[HLL_M];
  for (int i = 0; i < HLL_M; ++i) batch_regs[i] = (uint8_t)t->h_regs[i];
  const double batch_est = estimate ? 0.0 : hll_estimate_from_regs(batch_regs, hash_bits(t->hash_kind));   // used by insert_no_estimate only
  bool fits = true;
  if (estimate) {
    const double est = merge_batch_registers(t, batch_regs);
    reserve_for_estimate(t, est, t->size == 0);
    if (t->has_sticky) rebuild(t, t->capacity);                                       // the tiles carry no sticky bits
  } else {
    fits = (double)t->size + std::min((double)n, batch_est * (1.0 + hll_error_rate())) <= t->max_load * (double)t->capacity;
    if (!fits) {
      // The doubling schedule ends at the smallest capacity 128 * 2^j whose max_load bound holds the final size (the size only
      // grows).  The final size is at least `lo` (4 sigma of the batch estimate), so the schedule reaches capacity_for(lo) in
      // any case: grow there at once -- one rebuild instead of one per doubling and the general kernel in between -- provided
      // the batch cannot overfill that table (`hi`); a final doubling, if the size demands it, follows the insert as always.
      const double sig = 4.0 * hll_error_rate();
      const double lo = std::max((double)t->size, batch_est * (1.0 - sig));
      const double hi = (double)t->size + std::min((double)n, batch_est * (1.0 + sig));
      const uint64_t cap_lo = std::max(t->capacity, capacity_for(t, (uint64_t)lo));
      if (hi <= 0.5 * (t->max_load + 1.0) * (double)cap_lo) {
        if (cap_lo != t->capacity) rebuild(t, cap_lo, t->size == 0);
        fits = true;
      }
    }
    if (fits && t->has_sticky) rebuild(t, t->capacity);
  }
  const uint64_t tile_slots = (uint64_t)BK_TILE_BYTES / t->slot_bytes();
  const int lbk = log2_u64(std::max<uint64_t>(t->capacity / t->bucket_slots(), 1));
  const int tile_bits = lbk - BK_BUCKET_BITS;                        // log2(number of tiles)
  // ordered batches: a key on the overflow list has lost its batch position, the general route takes the whole batch
  const bool suits = fits && t->capacity >= tile_slots && n * 4 >= t->capacity && lbk <= 32 && tile_bits - b.lp <= BK_MAX_LP &&
                     n_ovf <= (b.ord ? 0ULL : b.ovf_cap);
  if (!suits) {
    b.release();
    DevBuf<uint64_t> flat;
    const uint64_t *keys = b.refetch ? b.refetch(flat, s) : flatten_segments(t, d_keys, n, segs, flat, s);
    insert_general(t, keys, b.vals, n, dflt, estimate != 0, s);
    KH_CUDA(cudaStreamSynchronize(s));
    return (int64_t)(t->size - size0);
  }
  const uint4 *recs = b.recs.p;
  const unsigned long long *ends = b.cursors.p;
  unsigned long long slab = b.slab;
  int rpt_log2 = b.lp - tile_bits;
  DevBuf<uint4> recs2;
  DevBuf<unsigned long long> l2buf;
  // more tiles than regions, at most 16 per region: a cluster of CTAs per region can insert into its tiles over distributed shared
  // memory, which saves the level-2 pass (kh_bulk.cuh: region_insert_kernel).  OFF by default: measured at 1 B keys / 16 tiles
  // per region the remote atomics cost 78 ms against 19.6 ms for level 2 + tile insert.  KH_CLUSTER=1 switches it on.
  const bool cluster_on = getenv("KH_CLUSTER") && atoi(getenv("KH_CLUSTER")) != 0;         // read per call: tests switch it
  const int cluster_tb = (cluster_on && !b.ord && tile_bits > b.lp && tile_bits - b.lp <= 4) ? tile_bits - b.lp : 0;
  if (tile_bits > b.lp && !cluster_tb) {
    // more tiles than regions: split every region by the next tag bits.  Default: into one SLAB per tile, sized for a
    // uniform hash like the level-1 slabs (no histogram pass; a record that does not fit hands its key to the overflow
    // list).  KH_L2_EXACT=1: histogram of the tags -> scan -> scatter into exactly packed runs.
    Trace tr("bulk.level2", s);
    const bool exact = getenv("KH_L2_EXACT") && atoi(getenv("KH_L2_EXACT")) != 0;      // read per call: tests switch it
    const int tbits = tile_bits - b.lp;
    const uint64_t R = 1ULL << b.lp, NT = 1ULL << tile_bits;
    const uint64_t slab2 = exact ? 0 : (((uint64_t)((double)std::max<uint64_t>(b.fed, 1) / (double)NT * 1.125) + 512 + 127) & ~127ULL);
    const bool fits32 = exact || NT * slab2 < (1ULL << 32);
    const bool use_slab = !exact && fits32;
    const int cstride = use_slab ? BK_CSTRIDE : 1;
    l2buf.alloc(R + 1 + (use_slab ? NT * cstride : 2 * NT));
    recs2.alloc((use_slab ? NT * slab2 : std::max<uint64_t>(b.fed, 1)) * W);
    L2Args a;
    a.recs = b.recs.p; a.cursors = b.cursors.p; a.slab = b.slab; a.lp = b.lp; a.tbits = tbits;
    a.chunk_prefix = l2buf.p; a.recs2 = recs2.p;
    a.slab2 = use_slab ? slab2 : 0; a.cstride = cstride;
    a.ovf_keys = b.ovf.p; a.ovf_count = b.ovf_count.p; a.ovf_cap = b.ovf_cap;
    KH_LAUNCH(l2_plan_kernel, 1, 1024, 0, s, a);
    const int grid = t->sms * 2;
    kev_begin(t, 5, s);
    DevBuf<unsigned char> tmp;
    if (use_slab) {
      a.tile_cnt = nullptr; a.tile_cur = l2buf.p + R + 1;
      KH_LAUNCH(l1_cursor_kernel, (int)((NT + 255) / 256), 256, 0, s, a.tile_cur, (int)NT, (unsigned long long)slab2);
    } else {
      a.tile_cnt = l2buf.p + R + 1; a.tile_cur = a.tile_cnt + NT;
      KH_CUDA(cudaMemsetAsync(a.tile_cnt, 0, NT * sizeof(unsigned long long), s));
      if (W == 1) KH_LAUNCH(l2_hist_kernel<1>, grid, L2_THREADS, 0, s, a); else KH_LAUNCH(l2_hist_kernel<2>, grid, L2_THREADS, 0, s, a);
      size_t tmp_bytes = 0;
      KH_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, a.tile_cnt, a.tile_cur, NT, s));
      tmp.alloc(tmp_bytes);
      KH_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, a.tile_cnt, a.tile_cur, NT, s));
      g_launches.fetch_add(1, std::memory_order_relaxed);
    }
    if (W == 1) {
      KH_CUDA(cudaFuncSetAttribute(l2_scatter_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)l2_smem_bytes(1)));
      KH_LAUNCH(l2_scatter_kernel<1>, grid, L2_THREADS, l2_smem_bytes(1), s, a);
    } else {
      KH_CUDA(cudaFuncSetAttribute(l2_scatter_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)l2_smem_bytes(2)));
      KH_LAUNCH(l2_scatter_kernel<2>, grid, L2_THREADS, l2_smem_bytes(2), s, a);
    }
    kev_end(t, s);
    if (use_slab) KH_CUDA(cudaMemcpyAsync(&n_ovf, b.ovf_count.p, sizeof(n_ovf), cudaMemcpyDeviceToHost, s));
    KH_CUDA(cudaStreamSynchronize(s));                               // tmp goes out of scope; the overflow count is final
    recs = recs2.p; ends = a.tile_cur; slab = use_slab ? slab2 : 0; rpt_log2 = 0;
    if (n_ovf > (b.ord ? 0ULL : b.ovf_cap)) {                        // a heavily skewed batch: the general route takes over
      b.release(); recs2.release();
      DevBuf<uint64_t> flat;
      const uint64_t *keys = b.refetch ? b.refetch(flat, s) : flatten_segments(t, d_keys, n, segs, flat, s);
      insert_general(t, keys, b.vals, n, dflt, estimate != 0, s);
      KH_CUDA(cudaStreamSynchronize(s));
      return (int64_t)(t->size - size0);
    }
  }
  uint32_t *const defer = b.defer.p;
  uint64_t n_def = 0;
  { Trace tr("bulk.tiles", s);
    zero_counter(t, s);
    TileArgs a;
    a.t = t->view(); a.recs = recs; a.ends = ends; a.slab = slab; a.rpt_log2 = rpt_log2; a.lbk = lbk;
    a.fresh = t->uninit ? 2 : (t->size == 0 ? 1 : 0); a.dflt = dflt; a.n_new = t->counters.p; a.defer_idx = defer;
    a.vals = b.vals;
    const int grid = (int)std::min<uint64_t>(t->capacity / tile_slots, (uint64_t)t->sms * 2);
    kev_begin(t, 0, s);
    if (b.ord) {
      const size_t smem = W == 1 ? tile_ord_smem_bytes<1>() : tile_ord_smem_bytes<2>();
      const void *fn = W == 1 ? (b.ord == INS_LAST ? (const void *)tile_insert_ord_kernel<1, true> : (const void *)tile_insert_ord_kernel<1, false>)
                              : (b.ord == INS_LAST ? (const void *)tile_insert_ord_kernel<2, true> : (const void *)tile_insert_ord_kernel<2, false>);
      KH_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      void *params[] = {(void *)&a};
      KH_CUDA(cudaLaunchKernel(fn, dim3((unsigned)grid), dim3(TI_THREADS), params, smem, s));
      g_launches.fetch_add(1, std::memory_order_relaxed);
    } else if (cluster_tb) {
      const int T = 1 << cluster_tb;
      const size_t smem = W == 1 ? tile_smem_bytes<1>() : tile_smem_bytes<2>();
      const void *fn = W == 1 ? (const void *)region_insert_kernel<1> : (const void *)region_insert_kernel<2>;
      KH_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      if (T > 8) KH_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      cudaLaunchConfig_t cfg{};
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = (unsigned)T; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
      cfg.blockDim = dim3(TI_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = s; cfg.attrs = attr; cfg.numAttrs = 1;
      cfg.gridDim = dim3((unsigned)T);
      int max_clusters = 0;
      KH_CUDA(cudaOccupancyMaxActiveClusters(&max_clusters, fn, &cfg));
      const uint64_t n_regions = (t->capacity / tile_slots) >> cluster_tb;
      const uint64_t ncl = std::max<uint64_t>(1, std::min<uint64_t>(n_regions, (uint64_t)std::max(max_clusters, 1)));
      cfg.gridDim = dim3((unsigned)(ncl * T));
      if (W == 1) KH_CUDA(cudaLaunchKernelEx(&cfg, region_insert_kernel<1>, a, cluster_tb));
      else KH_CUDA(cudaLaunchKernelEx(&cfg, region_insert_kernel<2>, a, cluster_tb));
      g_launches.fetch_add(1, std::memory_order_relaxed);
    } else if (W == 1) {
      KH_CUDA(cudaFuncSetAttribute(tile_insert_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_smem_bytes<1>()));
      KH_LAUNCH(tile_insert_kernel<1>, grid, TI_THREADS, tile_smem_bytes<1>(), s, a);
    } else {
      KH_CUDA(cudaFuncSetAttribute(tile_insert_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_smem_bytes<2>()));
      KH_LAUNCH(tile_insert_kernel<2>, grid, TI_THREADS, tile_smem_bytes<2>(), s, a);
    }
    kev_end(t, s);
    t->uninit = false;                                               // enqueued: every tile gets written
    if (!b.ord) {
      // chains that leave their tile, the all-ones key: a handful of keys through the global kernel, which takes them from the
      // records and their number from the tile kernel's counter (no host round trip in between); it probes the whole table
      const RecSource src{t->counters.p + 1, recs, defer};
      launch_insert(t, nullptr, nullptr, n, dflt, INS_COUNT, s, nullptr, false, &src);
    }
    t->size += read_counter(t, s);
    n_def = b.ord ? t->h_counters[1] : 0; }
  if (n_def) {       // ordered tables: the deferred keys in batch order, with their values
    Trace tr("bulk.deferred", s);
    DevBuf<uint64_t> nk(n_def * W);
    {                // sort the (position, record) pairs of the deferred keys by position
      DevBuf<uint32_t> pos(n_def), pos_sorted(n_def), idx_sorted(n_def), nv(b.vals ? n_def : 0);
      const int g = (int)((n_def + 255) / 256);
      if (W == 1) KH_LAUNCH((gather_rec_pos_kernel<1>), g, 256, 0, s, recs, defer, n_def, pos.p);
      else KH_LAUNCH((gather_rec_pos_kernel<2>), g, 256, 0, s, recs, defer, n_def, pos.p);
      size_t tmp_bytes = 0;
      KH_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, pos.p, pos_sorted.p, defer, idx_sorted.p, (int)n_def, 0, 32, s));
      DevBuf<unsigned char> tmp(tmp_bytes);
      KH_CUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, pos.p, pos_sorted.p, defer, idx_sorted.p, (int)n_def, 0, 32, s));
      g_launches.fetch_add(1, std::memory_order_relaxed);
      if (W == 1) KH_LAUNCH((gather_rec_pairs_kernel<1>), g, 256, 0, s, recs, idx_sorted.p, pos_sorted.p, b.vals, n_def, nk.p, nv.p);
      else KH_LAUNCH((gather_rec_pairs_kernel<2>), g, 256, 0, s, recs, idx_sorted.p, pos_sorted.p, b.vals, n_def, nk.p, nv.p);
      insert_resident(t, nk.p, b.vals ? nv.p : nullptr, n_def, dflt, s);
      KH_CUDA(cudaStreamSynchronize(s));                             // the temporaries go out of scope
    }
  }
  if (n_ovf) insert_resident(t, b.ovf.p, nullptr, n_ovf, dflt, s);
  KH_CUDA(cudaStreamSynchronize(s));
  return (int64_t)(t->size - size0);
}

int64_t insert_bulk(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n, uint32_t dflt, int estimate,
                    const BkSegs *segs, cudaStream_t s) {
  Trace tr_ins("insert.total(bulk)", s);
  struct Materialize {                  // whatever happens: never leave uninitialised slots behind
    kh_table *t; cudaStream_t s;
    ~Materialize() { if (t->uninit) { try { materialize(t, s); cudaStreamSynchronize(s); } catch (...) {} } }
  } guard{t, s};
  BulkBatch b;
  if (t->reducer != KH_REDUCE_COUNT) { b.ord = t->reducer == KH_REDUCE_FIRST ? INS_FIRST : INS_LAST; b.vals = d_vals; }
  bulk_begin(b, t, n, s);
  { Trace tr("bulk.level1", s); bulk_feed(b, t, d_keys, n, s, segs); }
  const int64_t got = bulk_finish(b, t, d_keys, segs, dflt, estimate, s);
  while ((double)t->size > t->max_load * (double)t->capacity) rebuild(t, t->capacity * 2);
  return got;
}

}  // namespace

int64_t kh::query_pieces_dev(kh_table *t, int op, int n_pieces, const uint64_t *const *d_keys, const uint64_t *n, void *const *d_out,
                             uint32_t nonexistent) {
  cudaStream_t s0 = t->stream[0], s1 = t->stream[1];
  zero_counter(t, s0);
  KH_CUDA(cudaEventRecord(t->ev[0], s0));
  KH_CUDA(cudaStreamWaitEvent(s1, t->ev[0], 0));
  int used = 0;
  for (int q = 0; q < n_pieces; ++q) {
    if (!n[q]) continue;
    cudaStream_t s = (used++ & 1) ? s1 : s0;
    if (op == 0) launch_lookup(t, OP_FIND, d_keys[q], n[q], reinterpret_cast<uint32_t *>(d_out[q]), nullptr, nullptr, nonexistent, t->counters.p, s);
    else launch_lookup(t, OP_COUNT, d_keys[q], n[q], nullptr, reinterpret_cast<uint8_t *>(d_out[q]), nullptr, 0, t->counters.p, s);
  }
  KH_CUDA(cudaEventRecord(t->ev[1], s1));
  KH_CUDA(cudaStreamWaitEvent(s0, t->ev[1], 0));
  return (int64_t)read_counter(t, s0);
}

// ---- the bulk insert in pieces, for the pipelined multi-GPU receiver (exchange.cu) -------------------------------------------
void kh::bulk_pipe_begin(BulkBatch &b, kh_table *t, uint64_t n_upper, cudaStream_t s) { bulk_begin(b, t, n_upper, s); }
void kh::bulk_pipe_feed(BulkBatch &b, kh_table *t, const uint64_t *d_base, uint64_t m_upper, const BkSegs *dsegs, cudaStream_t s) {
  bulk_feed(b, t, d_base, m_upper, s, nullptr, dsegs);
}
int64_t kh::bulk_pipe_finish(BulkBatch &b, kh_table *t, uint64_t n_recv, uint32_t dflt, cudaStream_t s) {
  struct Materialize {                  // whatever happens: never leave uninitialised slots behind
    kh_table *t; cudaStream_t s;
    ~Materialize() { if (t->uninit) { try { materialize(t, s); cudaStreamSynchronize(s); } catch (...) {} } }
  } guard{t, s};
  b.n = b.fed = n_recv;
  const int64_t got = bulk_finish(b, t, nullptr, nullptr, dflt, 1, s);
  while ((double)t->size > t->max_load * (double)t->capacity) rebuild(t, t->capacity * 2);
  return got;
}
bool kh::bulk_pipe_candidate(const kh_table *t, uint64_t n) { return bulk_candidate(t, nullptr, n); }

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
    t->side.alloc(4); t->counters.alloc(3); t->status.p = reinterpret_cast<int *>(t->counters.p + 2);
    KH_CUDA(cudaMemset(t->side.p, 0, 4 * sizeof(uint32_t)));
    KH_CUDA(cudaHostAlloc((void **)&t->h_counters, 3 * sizeof(unsigned long long), cudaHostAllocDefault));
    t->h_status = reinterpret_cast<int *>(t->h_counters + 2);
    KH_CUDA(cudaHostAlloc((void **)&t->h_regs, (HLL_M + 2) * sizeof(uint32_t), cudaHostAllocDefault));
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
  if (t->h_regs) cudaFreeHost(t->h_regs);
  t->slots.release(); t->side.release(); t->counters.release();
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
  for (int i = 0; i < 6; ++i) { t->kev_ms[i] = 0.0; t->kev_n[i] = 0; }
  return KH_OK;
}
int kh_kernel_time(kh_table *t, int which, double *sum_ms, uint64_t *launches) {
  if (!t || which < 0 || which > 5) return KH_ERR_INVALID;
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
// estimate != 0 = the reference's insert(): HyperLogLog over the batch joins the table's registers, reserve for the estimate,
// then one pass (KH/robinhood_offset_hashmap_ptr.hpp:2512-2535).  estimate == 0 = insert_no_estimate(): fill up to max_load,
// double, continue (:2802-2885).  pre: batch registers the caller's pipeline already computed, and/or the pieces of the batch.
int64_t kh::insert_dev_impl(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n,
                            uint32_t dflt, int estimate, const PreEstimate *pre) {
  cudaStream_t s = t->stream[0];
  if (n == 0) return 0;
  const uint64_t size0 = t->size;
  const BkSegs *segs = pre ? pre->segs : nullptr;
  if (bulk_candidate(t, d_vals, n, segs)) return insert_bulk(t, d_keys, d_vals, n, dflt, estimate, segs, s);
  DevBuf<uint64_t> flat;
  d_keys = flatten_segments(t, d_keys, n, segs, flat, s);
  struct Materialize {                  // whatever path follows (or throws): never leave uninitialised slots behind
    kh_table *t; cudaStream_t s;
    ~Materialize() { if (t->uninit) { try { materialize(t, s); cudaStreamSynchronize(s); } catch (...) {} } }
  } mat_guard{t, s};
  const bool count_mode = t->reducer == KH_REDUCE_COUNT;
  bool single = false;
  if (estimate || (count_mode && n >= (1ULL << 16) && (double)(t->size + n) > t->max_load * (double)t->capacity)) {
    // insert_no_estimate on a table the caller reserved (the drivers: estimate -> reserve -> insert_no_estimate): the
    // HLL pass costs one streaming read and saves the piecewise schedule
    uint8_t regs[HLL_M];
    const uint8_t *batch_regs = pre ? pre->regs : nullptr;
    if (!batch_regs) {
      Trace tr("hll", s);
      kev_begin(t, 3, s);
      hll_estimate_dev(t->hash_kind, t->hp, t->W, d_keys, n, regs, s, t->sms);
      kev_end(t, s);
      batch_regs = regs;
    }
    if (estimate) {
      reserve_for_estimate(t, merge_batch_registers(t, batch_regs), false);
      single = count_mode;              // COUNT tolerates a wrong bound: keys that find no slot are deferred and re-inserted after growing
    } else {
      const double est = hll_estimate_from_regs(batch_regs, hash_bits(t->hash_kind));
      single = (double)t->size + std::min((double)n, est * (1.0 + hll_error_rate())) <= t->max_load * (double)t->capacity;
    }
  }
  Trace tr_ins("insert.total", s);
  insert_general(t, d_keys, d_vals, n, dflt, single, s);
  while ((double)t->size > t->max_load * (double)t->capacity) rebuild(t, t->capacity * 2);
  return (int64_t)(t->size - size0);
}

extern "C" {

int64_t kh_insert_dev(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n,
                      uint32_t default_val, int estimate) {
  return guarded([&]() -> int64_t {
    if (!t || (!d_keys && n)) { set_error("null table or keys"); return KH_ERR_INVALID; }
    DeviceGuard g(t->device);
    int64_t added = 0;
    for (uint64_t off = 0; off < n; off += BULK_MAX_KEYS) {          // record positions of the bulk route are 32-bit
      const uint64_t m = std::min(BULK_MAX_KEYS, n - off);
      const int64_t got = insert_dev_impl(t, d_keys + off * t->W, d_vals ? d_vals + off : nullptr, m, default_val, estimate);
      if (got < 0) return got;
      added += got;
    }
    return added;
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
    if (estimate && !vals && bulk_candidate(t, nullptr, n)) {
      // Bulk route, fed chunk by chunk BEHIND the PCIe copy: when the last chunk lands, the batch is already partitioned by
      // table region and its HLL registers are complete; only the estimate, the reserve and the tile insert remain.
      DevBuf<uint64_t> dk(n * W);
      BulkBatch b;
      bulk_begin(b, t, n, t->stream[0]);
      int q = 0;
      for (uint64_t off = 0; off < n; off += HOST_CHUNK, q ^= 1) {
        const uint64_t m = std::min(HOST_CHUNK, n - off);
        KH_CUDA(cudaMemcpyAsync(dk.p + off * W, keys + off * W, m * W * 8, cudaMemcpyHostToDevice, t->stream[1]));
        KH_CUDA(cudaEventRecord(t->ev[q], t->stream[1]));
        KH_CUDA(cudaStreamWaitEvent(t->stream[0], t->ev[q], 0));
        bulk_feed(b, t, dk.p + off * W, m, t->stream[0]);
      }
      KH_CUDA(cudaStreamSynchronize(t->stream[1]));
      struct Materialize {
        kh_table *t; cudaStream_t s;
        ~Materialize() { if (t->uninit) { try { materialize(t, s); cudaStreamSynchronize(s); } catch (...) {} } }
      } guard{t, t->stream[0]};
      const int64_t got = bulk_finish(b, t, dk.p, nullptr, default_val, estimate, t->stream[0]);
      while ((double)t->size > t->max_load * (double)t->capacity) rebuild(t, t->capacity * 2);
      return got;
    }
    if (estimate || t->reducer != KH_REDUCE_COUNT) {
      // Whole batch resident (the ordered reducers need stable batch positions, and presizing needs the
      // estimate before the first insert).  The HLL registers are accumulated chunk by chunk BEHIND the
      // PCIe copy, so the estimate is ready when the last chunk lands.
      DevBuf<uint64_t> dk(n * W);
      DevBuf<uint32_t> dv(vals ? n : 0);
      const bool pipe_hll = estimate != 0;
      DevBuf<uint32_t> regs(pipe_hll ? HLL_M : 0);
      if (pipe_hll) hll_reset(regs.p, nullptr, t->stream[0]);
      int b = 0;
      for (uint64_t off = 0; off < n; off += HOST_CHUNK, b ^= 1) {
        const uint64_t m = std::min(HOST_CHUNK, n - off);
        KH_CUDA(cudaMemcpyAsync(dk.p + off * W, keys + off * W, m * W * 8, cudaMemcpyHostToDevice, t->stream[1]));
        if (vals) KH_CUDA(cudaMemcpyAsync(dv.p + off, vals + off, m * 4, cudaMemcpyHostToDevice, t->stream[1]));
        if (pipe_hll) {
          KH_CUDA(cudaEventRecord(t->ev[b], t->stream[1]));
          KH_CUDA(cudaStreamWaitEvent(t->stream[0], t->ev[b], 0));
          hll_accumulate(t->hash_kind, t->hp, W, dk.p + off * W, m, regs.p, nullptr, t->stream[0], t->sms);
        }
      }
      KH_CUDA(cudaStreamSynchronize(t->stream[1]));
      if (!pipe_hll) return insert_dev_impl(t, dk.p, vals ? dv.p : nullptr, n, default_val, 0);
      uint8_t batch_regs[HLL_M];
      hll_finish(t->hash_kind, regs.p, batch_regs, t->stream[0]);
      PreEstimate pre{batch_regs, nullptr};
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
