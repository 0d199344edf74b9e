// kh_hll.cuh -- HyperLogLog distinct estimate on the GPU (precision 12, 4096 registers).
//
// Replaces hyperloglog64<Key, Hash, 12>::update + estimate (KH/hyperloglog64.hpp:145-239) as
// used by insert_impl (KH/robinhood_offset_hashmap_ptr.hpp:2512-2535) and the driver's
// "estimate" phase (benchmark/BenchmarkHashTables.cpp:1232-1248).  Register contents are
// bit-identical to the reference's for the same keys and store hash:
//   idx  = hash >> (bits - 12)
//   rank = clz((hash << 12) | 0xFFF) + 1           (in the hash's own width, 64 or 32)
// Per block: 4096 x u32 shared-memory registers with atomicMax, merged into global once.
#pragma once
#include <cstdint>
#include <cmath>
#include "kh_common.h"
#include "kh_table.cuh"

namespace kh {

constexpr int HLL_P = 12;
constexpr int HLL_M = 1 << HLL_P;

inline double hll_error_rate() { return 1.04 / std::sqrt((double)HLL_M); }

// number of fine table regions the fused pass histograms (top REGION_FINE_BITS bits of the spread hash)
constexpr int REGION_FINE_BITS = 10;
constexpr int REGION_FINE = 1 << REGION_FINE_BITS;

// g_hist != nullptr: the same pass also counts the keys per fine table region (the windowed bulk insert
// derives its partition offsets from it, so the keys are hashed and read once for both purposes).
template <int W>
__global__ void __launch_bounds__(512) hll_update_kernel(const uint64_t *keys, uint64_t n, int hash_kind, HashParams hp,
                                                         uint32_t *g_regs, unsigned long long *g_hist) {
  __shared__ uint32_t regs[HLL_M];
  __shared__ uint32_t fine[REGION_FINE];
  for (int i = threadIdx.x; i < HLL_M; i += blockDim.x) regs[i] = 0;
  for (int i = threadIdx.x; i < REGION_FINE; i += blockDim.x) fine[i] = 0;
  __syncthreads();
  const bool h32 = hash_bits(hash_kind) == 32;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t key[W];
    load_key<W>(keys, i, key);
    const uint64_t h = hash_runtime<W>(hash_kind, key, hp);
    uint32_t idx, rank;
    if (h32) {
      const uint32_t x = (uint32_t)h;
      idx = x >> (32 - HLL_P);
      rank = __clz((int)((x << HLL_P) | 0xFFFu)) + 1;
    } else {
      idx = (uint32_t)(h >> (64 - HLL_P));
      rank = __clzll((long long)((h << HLL_P) | 0xFFFULL)) + 1;
    }
    if (regs[idx] < rank) atomicMax(&regs[idx], rank);
    if (g_hist) atomicAdd(&fine[(uint32_t)(spread_hash(h) >> (64 - REGION_FINE_BITS))], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < HLL_M; i += blockDim.x)
    if (regs[i]) atomicMax(&g_regs[i], regs[i]);
  if (g_hist)
    for (int i = threadIdx.x; i < REGION_FINE; i += blockDim.x)
      if (fine[i]) atomicAdd(&g_hist[i], (unsigned long long)fine[i]);
}

// estimate() of the reference (hyperloglog64.hpp:203-239): raw harmonic estimate, linear counting
// for small cardinalities, large-range correction only for 32-bit hashes.
inline double hll_estimate_from_regs(const uint8_t *regs, int bits) {
  const double m = (double)HLL_M;
  double amm = 0.7213 / (1.0 + 1.079 / m);
  amm *= m * m;
  // 2^-r from a table: the same exact powers of two, added in the same order, as ldexp gives -- at a tenth of the time (this runs
  // on the host between two kernels of every insert batch, with the GPU waiting)
  static const struct Pow2Table { double v[256]; Pow2Table() { for (int r = 0; r < 256; ++r) v[r] = std::ldexp(1.0, -r); } } pow2;
  double sum = 0.0;
  uint32_t zeros = 0;
  for (int i = 0; i < HLL_M; ++i) { sum += pow2.v[regs[i]]; zeros += regs[i] == 0; }
  double est = amm / sum;
  if (est <= 5.0 * (m / 2.0)) {
    if (zeros > 0) return m * std::log(m / (double)zeros);
    return est;
  }
  if (bits == 32 && est > (double)(1ULL << 31) / 15.0) return -std::log(1.0 - est / 4294967296.0) * 4294967296.0;
  return est;
}

// Accumulating form: registers (HLL_M x u32) and the optional fine region histogram live on the device
// and are updated batch by batch (the host pipeline feeds chunks as they arrive over PCIe).
inline void hll_reset(uint32_t *d_regs, unsigned long long *d_hist, cudaStream_t s) {
  KH_CUDA(cudaMemsetAsync(d_regs, 0, HLL_M * sizeof(uint32_t), s));
  if (d_hist) KH_CUDA(cudaMemsetAsync(d_hist, 0, REGION_FINE * sizeof(unsigned long long), s));
}
inline void hll_accumulate(int hash_kind, const HashParams &hp, int W, const uint64_t *d_keys, uint64_t n,
                           uint32_t *d_regs, unsigned long long *d_hist, cudaStream_t s, int sms) {
  if (!n) return;
  const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 2047) / 2048, (uint64_t)sms * 4));
  if (W == 1) KH_LAUNCH((hll_update_kernel<1>), grid, 512, 0, s, d_keys, n, hash_kind, hp, d_regs, d_hist);
  else KH_LAUNCH((hll_update_kernel<2>), grid, 512, 0, s, d_keys, n, hash_kind, hp, d_regs, d_hist);
}
// reads the registers back (synchronises s) and evaluates the estimate
inline double hll_finish(int hash_kind, const uint32_t *d_regs, uint8_t *regs_out, cudaStream_t s) {
  uint32_t h[HLL_M];
  KH_CUDA(cudaMemcpyAsync(h, d_regs, sizeof(h), cudaMemcpyDeviceToHost, s));
  KH_CUDA(cudaStreamSynchronize(s));
  uint8_t r8[HLL_M];
  for (int i = 0; i < HLL_M; ++i) r8[i] = (uint8_t)h[i];
  if (regs_out) for (int i = 0; i < HLL_M; ++i) regs_out[i] = r8[i];
  return hll_estimate_from_regs(r8, hash_bits(hash_kind));
}

// d_hist (device, REGION_FINE entries, may be null): zeroed and filled with the fine region histogram
inline double hll_estimate_dev(int hash_kind, const HashParams &hp, int W, const uint64_t *d_keys, uint64_t n,
                               uint8_t *regs_out, cudaStream_t s, int sms, unsigned long long *d_hist = nullptr) {
  DevBuf<uint32_t> g(HLL_M);
  hll_reset(g.p, d_hist, s);
  hll_accumulate(hash_kind, hp, W, d_keys, n, g.p, d_hist, s, sms);
  return hll_finish(hash_kind, g.p, regs_out, s);
}

}  // namespace kh
