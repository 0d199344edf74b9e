// kh_partition.h -- interface between partition.cu and dist.cu
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <vector>
#include "kh_hash.cuh"
#include "kh_common.h"

namespace kh {

// destination segment r starts at base[r] (device pointers, possibly on peer GPUs); keys land at
// base[r] + (global position - seg_start[r]) * W.
struct PartDst {
  uint64_t *base[8];
};

struct PartitionPlan {
  int device = 0, kind = 0, W = 1;
  HashParams hp;
  uint32_t p = 1;
  int region_shift = 0;          // > 0: partition by table region (top bits of the spread store hash), p = 2^(64-shift)
  const uint64_t *d_keys = nullptr;
  uint64_t n = 0, blocks = 0, per_block = 0;
  DevBuf<uint64_t> hist, offs, seg;
  std::vector<uint64_t> seg_h;   // [p+1] start of every destination segment in the permuted order
  void init(int device, int kind, const HashParams &hp, int W, uint32_t p, const uint64_t *d_keys, uint64_t n);
  void count(cudaStream_t s, uint64_t *counts_host);          // synchronises s
  void scatter(cudaStream_t s, uint64_t *d_out_keys, uint32_t *d_out_perm, const PartDst *dst);   // asynchronous
};

// counts_host[p] receives the per-rank totals.  When dst != nullptr the keys are written through
// dst->base[] (p <= 8) instead of d_out_keys.  Synchronises `s` before returning.
void partition_dev(int device, int kind, const HashParams &hp, int W, uint32_t p, const uint64_t *d_keys, uint64_t n,
                   uint64_t *d_out_keys, uint32_t *d_out_perm, uint64_t *counts_host, const PartDst *dst,
                   cudaStream_t s);

}  // namespace kh
