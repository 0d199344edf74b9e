// kh_table_host.h -- the table object and the internal entry points shared by table.cu and exchange.cu
#pragma once
#include <algorithm>
#include <functional>
#include <utility>
#include <vector>
#include "kh_common.h"
#include "kh_table.cuh"
#include "kh_hll.cuh"

using namespace kh;   // internal header: only table.cu and exchange.cu include it

struct kh_table {
  int device = 0, map_kind = 0, W = 1, hash_kind = 0, reducer = 0;
  uint32_t seed = KH_DEFAULT_SEED;
  double min_load = 0.2, max_load = 0.8;
  int LB = 32, G = 2;                 // probe geometry (DESIGN.md "bucket geometry")
  bool window_on = true;              // L2-windowed bulk insert (KH_WINDOW=0 switches it off for A/B runs)
  uint64_t window_bytes = 32ULL << 20, window_min_keys = 1ULL << 22;
  bool uninit = false;                // slots allocated but not yet initialised (insert_dev_impl -> tile insert writes every tile)
  bool has_sticky = false;            // an erase may have left sticky bits since the last rebuild / clear (the tile insert needs none)
  uint8_t hll[HLL_M] = {0};           // HyperLogLog registers over every batch inserted with an estimate (the reference's
                                      // cumulative this->hll, KH/robinhood_offset_hashmap_ptr.hpp:291,2520-2535)
  uint64_t capacity = 0, size = 0;
  HashParams hp;
  DevBuf<unsigned char> slots;
  DevBuf<uint32_t> side;              // 3 words
  DevBuf<unsigned long long> counters;  // [0] n_new / n_hit, [1] export counter, [2] = the status word below
  struct { int *p = nullptr; } status;  // lives in counters[2]: one memset clears, one copy reads counters and status
  unsigned long long *h_counters = nullptr;  // pinned mirror (3 words)
  int *h_status = nullptr;              // = h_counters + 2
  uint32_t *h_regs = nullptr;           // pinned: HLL_M registers + the overflow count of a bulk batch (one copy, no staging)
  cudaStream_t stream[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  // staging for *_host calls (grown on demand)
  DevBuf<uint64_t> d_keys[2];
  DevBuf<uint32_t> d_vals[2];
  DevBuf<uint8_t> d_u8[2];
  int sms = 148;
  // CUDA-event pairs around every probe-kernel launch (bench.py's roofline: measured on the launching stream)
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> kev;
  size_t kev_used = 0;
  // accumulated per kind: 0 table insert (tile insert, global insert kernel), 1 lookup kernels, 2 level-1 scatter, 3 stand-alone HLL,
  // 4 rebuild (every kernel launched inside rebuild(): streaming rebuild, dump, re-insert), 5 level-2 split (histogram, scan, scatter)
  double kev_ms[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  uint64_t kev_n[6] = {0, 0, 0, 0, 0, 0};
  bool kev_in_rebuild = false;
  std::vector<int> kev_kind;
  bool kev_on = false;

  int slot_bytes() const { return 16 * W; }
  int bucket_slots() const { return G * (LB / slot_bytes()); }
  TableView view() { return TableView{slots.p, capacity, side.p}; }
  int grid_for(uint64_t n_items, int per_block) const {
    uint64_t need = (n_items + per_block - 1) / per_block;
    uint64_t cap = (uint64_t)sms * 8;
    return (int)std::max<uint64_t>(1, std::min(need, cap));
  }
};


namespace kh {
// state of one bulk-insert batch (table.cu): begin -> feed (once, or chunk by chunk as the keys arrive over PCIe or NVLink) -> finish
struct BulkBatch {
  uint64_t n = 0, fed = 0, slab = 0, ovf_cap = 0;
  int lp = 0;
  // ONE allocation (every allocation costs a round trip of ~10 us during which the GPU idles), carved into the arrays below
  template <typename T> struct Part { T *p = nullptr; };
  DevBuf<unsigned char> arena;
  Part<unsigned long long> cursors, ovf_count;
  Part<uint4> recs;
  Part<uint64_t> ovf;
  Part<uint32_t> regs, defer;                  // defer: record positions of the keys the tile insert hands on (n entries)
  void release() { arena.release(); cursors.p = ovf_count.p = nullptr; recs.p = nullptr; ovf.p = nullptr; regs.p = defer.p = nullptr; }
  // FIRST / LAST tables: the records carry batch positions (ord = INS_FIRST / INS_LAST), vals = the batch's values (may be null)
  int ord = 0;
  const uint32_t *vals = nullptr;
  // the whole batch again as ONE array, for the rare case that the general route has to take over (table that does not suit the
  // tiles, skewed batch); unset: the caller's key pointer / pieces are used
  std::function<const uint64_t *(DevBuf<uint64_t> &, cudaStream_t)> refetch;
};
struct BkSegs;
// the bulk insert in pieces (pipelined multi-GPU receiver): n_upper / m_upper bound the keys, dsegs = DEVICE-resident pieces
void bulk_pipe_begin(BulkBatch &b, kh_table *t, uint64_t n_upper, cudaStream_t s);
void bulk_pipe_feed(BulkBatch &b, kh_table *t, const uint64_t *d_base, uint64_t m_upper, const BkSegs *dsegs, cudaStream_t s);
int64_t bulk_pipe_finish(BulkBatch &b, kh_table *t, uint64_t n_recv, uint32_t dflt, cudaStream_t s);
bool bulk_pipe_candidate(const kh_table *t, uint64_t n);
// find (op 0: 4-byte answers) / count (op 1: 1-byte answers) over several key arrays with one answer array each (the per-sender
// slabs of a peer receive buffer, exchange.cu): the launches go out back to back on the table's two streams, ONE counter
// read-back and synchronisation at the end.  Returns the number of hits.
int64_t query_pieces_dev(kh_table *t, int op, int n_pieces, const uint64_t *const *d_keys, const uint64_t *n, void *const *d_out,
                         uint32_t nonexistent);
// what a caller's pipeline already knows about a batch: its HyperLogLog registers (HLL_M bytes on the HOST, may be null:
// accumulated behind the PCIe copy) and/or the pieces it lies in (kh_bulk.cuh: BkSegs; the slabs of a peer receive buffer)
struct PreEstimate { const uint8_t *regs; const BkSegs *segs; };
// batch insert of device-resident keys (see table.cu); returns #new keys or a negative status
int64_t insert_dev_impl(kh_table *t, const uint64_t *d_keys, const uint32_t *d_vals, uint64_t n, uint32_t dflt,
                        int estimate, const PreEstimate *pre = nullptr);
}  // namespace kh
