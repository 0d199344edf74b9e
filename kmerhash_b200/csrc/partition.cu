// partition.cu -- hash partitioning of a key batch by owner rank (the step in front of the
// multi-GPU exchange) + its C ABI.
//
// Replaces:
//   modulus / key_to_hash         KH/distributed_batched_robinhood_map.hpp:164-178,182,652,749
//   assign_count_permute          :632-741   rank id per key + per-rank counts
//   permute_by_bucketid           :291-380   stable scatter into p contiguous segments
// rank = DistHash_{seed}(key) & (p-1) when p is a power of two, % p otherwise, evaluated in the
// functor's own result width.  The reference makes two scalar passes with a random scatter; here
// it is histogram (shared-memory atomics) -> one small scan -> stable scatter where a warp ranks
// its 32 keys with __match_any_sync and the 8 warps of a block are ordered through shared memory.
// The scatter can write each destination segment through its own base pointer, which is how
// dist.cu stores straight into the peer GPUs' receive buffers over NVLink.
#include <algorithm>
#include <mutex>
#include <vector>
#include <cub/device/device_scan.cuh>
#include "kh_common.h"
#include "kh_table.cuh"
#include "kh_partition.h"
#include "kh_window.cuh"

namespace kh {

constexpr int PART_MAX_P = 1024;
constexpr int PART_BLOCK = 256;
constexpr int PART_WARPS = PART_BLOCK / 32;

// region_shift == 0: owner rank of the exchange (the reference's rule).  region_shift > 0: table region =
// top bits of the spread store hash (kh_table.cuh), used by the L2-windowed bulk insert.
__device__ __forceinline__ uint32_t rank_of(uint64_t h, uint32_t p, bool pow2, int region_shift = 0) {
  if (region_shift) return (uint32_t)top_bits(spread_hash(h), region_shift);
  return pow2 ? (uint32_t)(h & (uint64_t)(p - 1)) : (uint32_t)(h % (uint64_t)p);
}

template <int W>
__global__ void rank_kernel(const uint64_t *keys, uint64_t n, int kind, HashParams hp, uint32_t p, uint32_t *out) {
  const bool pow2 = (p & (p - 1)) == 0;
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t key[W];
    load_key<W>(keys, i, key);
    out[i] = rank_of(hash_runtime<W>(kind, key, hp), p, pow2);
  }
}

// block b owns keys [b * per_block, (b+1) * per_block); hist layout [rank][block]
template <int W>
__global__ void __launch_bounds__(PART_BLOCK) part_hist_kernel(const uint64_t *keys, uint64_t n, uint64_t per_block,
                                                               int kind, HashParams hp, uint32_t p, int region_shift,
                                                               uint64_t *hist) {
  __shared__ uint32_t sh[PART_MAX_P];
  for (uint32_t r = threadIdx.x; r < p; r += blockDim.x) sh[r] = 0;
  __syncthreads();
  const bool pow2 = (p & (p - 1)) == 0;
  const uint64_t beg = (uint64_t)blockIdx.x * per_block;
  const uint64_t end = min(beg + per_block, n);
  for (uint64_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
    uint64_t key[W];
    load_key<W>(keys, i, key);
    const uint32_t r = rank_of(hash_runtime<W>(kind, key, hp), p, pow2, region_shift);
    // warp-aggregated shared atomic
    const uint32_t m = __match_any_sync(__activemask(), r);
    if ((int)(__ffs(m) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&sh[r], (uint32_t)__popc(m));
  }
  __syncthreads();
  for (uint32_t r = threadIdx.x; r < p; r += blockDim.x) hist[(uint64_t)r * gridDim.x + blockIdx.x] = sh[r];
}


template <int W, bool MULTI_DST>
__global__ void __launch_bounds__(PART_BLOCK) part_scatter_kernel(const uint64_t *keys, uint64_t n, uint64_t per_block,
                                                                  int kind, HashParams hp, uint32_t p, int region_shift,
                                                                  const uint64_t *offsets /*[rank][block] exclusive*/,
                                                                  uint64_t *out_keys, uint32_t *out_perm, PartDst dst,
                                                                  const uint64_t *seg_start /*[p+1], MULTI_DST only*/) {
  __shared__ uint64_t run[PART_MAX_P];
  __shared__ uint32_t wc[PART_WARPS][PART_MAX_P];
  for (uint32_t r = threadIdx.x; r < p; r += blockDim.x) {
    run[r] = offsets[(uint64_t)r * gridDim.x + blockIdx.x];
#pragma unroll
    for (int w = 0; w < PART_WARPS; ++w) wc[w][r] = 0;
  }
  __syncthreads();
  const bool pow2 = (p & (p - 1)) == 0;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt = (1u << lane) - 1u;
  const uint64_t beg = (uint64_t)blockIdx.x * per_block;
  const uint64_t end = min(beg + per_block, n);
  for (uint64_t c0 = beg; c0 < end; c0 += PART_BLOCK) {     // uniform trip count for the whole block
    const uint64_t i = c0 + threadIdx.x;
    const bool valid = i < end;
    uint64_t key[W];
#pragma unroll
    for (int q = 0; q < W; ++q) key[q] = 0;
    uint32_t r = p;                                          // invalid lanes form their own group
    if (valid) { load_key<W>(keys, i, key); r = rank_of(hash_runtime<W>(kind, key, hp), p, pow2, region_shift); }
    const uint32_t m = __match_any_sync(FULL_MASK, r);
    const bool leader = valid && ((__ffs(m) - 1) == lane);
    const uint32_t in_warp = __popc(m & lt);
    if (leader) wc[warp][r] = __popc(m);
    __syncthreads();
    uint64_t pos = 0;
    if (valid) {
      uint32_t before = 0;
#pragma unroll
      for (int w = 0; w < PART_WARPS; ++w) before += (w < warp) ? wc[w][r] : 0u;
      pos = run[r] + before + in_warp;
    }
    __syncthreads();
    if (leader) { atomicAdd((unsigned long long *)&run[r], (unsigned long long)__popc(m)); wc[warp][r] = 0; }
    if (valid) {
      uint64_t *o;
      if (MULTI_DST) o = dst.base[r] + (pos - seg_start[r]) * W;
      else o = out_keys + pos * W;
#pragma unroll
      for (int q = 0; q < W; ++q) o[q] = key[q];
      if (out_perm) out_perm[pos] = (uint32_t)i;
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void unpermute_kernel(const T *in, const uint32_t *perm, uint64_t n, T *out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[perm[i]] = in[i];
}

// ---- PartitionPlan: count() = histogram + scan (per-rank totals known on the host),
//      scatter() = stable scatter, optionally through per-destination base pointers -----------
void PartitionPlan::init(int device_, int kind_, const HashParams &hp_, int W_, uint32_t p_, const uint64_t *d_keys_,
                         uint64_t n_) {
  device = device_; kind = kind_; hp = hp_; W = W_; p = p_; d_keys = d_keys_; n = n_;
  seg_h.assign(p + 1, 0);
  blocks = per_block = 0;
  if (n == 0) return;
  const uint64_t max_blocks = (uint64_t)sm_count(device) * 8;
  blocks = std::min<uint64_t>(max_blocks, (n + PART_BLOCK - 1) / PART_BLOCK);
  per_block = (n + blocks - 1) / blocks;
  per_block = (per_block + PART_BLOCK - 1) / PART_BLOCK * PART_BLOCK;
  blocks = (n + per_block - 1) / per_block;
}

void PartitionPlan::count(cudaStream_t s, uint64_t *counts_host) {
  for (uint32_t r = 0; r < p; ++r) counts_host[r] = 0;
  if (n == 0) return;
  const uint64_t cells = (uint64_t)p * blocks;
  hist.alloc(cells + 1); offs.alloc(cells + 1); seg.alloc(p + 1);
  KH_CUDA(cudaMemsetAsync(hist.p + cells, 0, sizeof(uint64_t), s));
  if (W == 1) KH_LAUNCH((part_hist_kernel<1>), (int)blocks, PART_BLOCK, 0, s, d_keys, n, per_block, kind, hp, p, region_shift, hist.p);
  else KH_LAUNCH((part_hist_kernel<2>), (int)blocks, PART_BLOCK, 0, s, d_keys, n, per_block, kind, hp, p, region_shift, hist.p);
  size_t tmp_bytes = 0;
  KH_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, hist.p, offs.p, cells + 1, s));
  DevBuf<unsigned char> tmp(tmp_bytes);
  KH_CUDA(cub::DeviceScan::ExclusiveSum(tmp.p, tmp_bytes, hist.p, offs.p, cells + 1, s));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  // segment start of rank r = offsets[r][block 0]
  KH_CUDA(cudaMemcpy2DAsync(seg_h.data(), sizeof(uint64_t), offs.p, blocks * sizeof(uint64_t), sizeof(uint64_t), p,
                            cudaMemcpyDeviceToHost, s));
  KH_CUDA(cudaStreamSynchronize(s));   // also keeps `tmp` alive until the scan is done
  seg_h[p] = n;
  for (uint32_t r = 0; r < p; ++r) counts_host[r] = seg_h[r + 1] - seg_h[r];
  KH_CUDA(cudaMemcpyAsync(seg.p, seg_h.data(), (p + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, s));
}

void PartitionPlan::scatter(cudaStream_t s, uint64_t *d_out_keys, uint32_t *d_out_perm, const PartDst *dst) {
  if (n == 0) return;
  PartDst d{};
  if (dst) {
    d = *dst;
    if (W == 1) KH_LAUNCH((part_scatter_kernel<1, true>), (int)blocks, PART_BLOCK, 0, s, d_keys, n, per_block, kind, hp, p, region_shift, offs.p, d_out_keys, d_out_perm, d, seg.p);
    else KH_LAUNCH((part_scatter_kernel<2, true>), (int)blocks, PART_BLOCK, 0, s, d_keys, n, per_block, kind, hp, p, region_shift, offs.p, d_out_keys, d_out_perm, d, seg.p);
  } else {
    if (W == 1) KH_LAUNCH((part_scatter_kernel<1, false>), (int)blocks, PART_BLOCK, 0, s, d_keys, n, per_block, kind, hp, p, region_shift, offs.p, d_out_keys, d_out_perm, d, nullptr);
    else KH_LAUNCH((part_scatter_kernel<2, false>), (int)blocks, PART_BLOCK, 0, s, d_keys, n, per_block, kind, hp, p, region_shift, offs.p, d_out_keys, d_out_perm, d, nullptr);
  }
}

void partition_dev(int device, int kind, const HashParams &hp, int W, uint32_t p, const uint64_t *d_keys, uint64_t n,
                   uint64_t *d_out_keys, uint32_t *d_out_perm, uint64_t *counts_host, const PartDst *dst,
                   cudaStream_t s) {
  PartitionPlan plan;
  plan.init(device, kind, hp, W, p, d_keys, n);
  plan.count(s, counts_host);
  plan.scatter(s, d_out_keys, d_out_perm, dst);
  KH_CUDA(cudaStreamSynchronize(s));
}

// ---- static scratch of kh_partition_fast_noalloc_dev ---------------------------------------------------------------------------------
// hist | cursors | totals, KH_PARTITION_NOALLOC_MAX_P counters each.  In host code the array's NAME is a host-side stand-in, not a
// device address: the address on a given device comes from cudaGetSymbolAddress with that device current (once per device).
static __device__ __align__(256) unsigned long long g_part_scratch[3 * KH_PARTITION_NOALLOC_MAX_P];

static unsigned long long *noalloc_scratch(int device) {            // the caller has made `device` current
  static std::mutex mu;
  static unsigned long long *addr[64] = {nullptr};
  if (device < 0 || device >= 64) { set_error("device %d out of range", device); throw CudaError{KH_ERR_INVALID}; }
  std::lock_guard<std::mutex> lk(mu);
  if (!addr[device]) {
    unsigned long long *ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_part_scratch);
    if (e != cudaSuccess || !ptr) {
      cudaGetLastError();
      set_error("cudaGetSymbolAddress(g_part_scratch) on device %d: %s", device, cudaGetErrorString(e));
      throw CudaError{KH_ERR_CUDA};
    }
    addr[device] = ptr;
  }
  return addr[device];
}

}  // namespace kh

using namespace kh;

extern "C" {

int kh_partition_dev(int device, int hash_kind, uint32_t seed, int key_words, uint32_t p,
                     const uint64_t *d_keys, uint64_t n, uint64_t *d_out_keys, uint32_t *d_out_perm,
                     uint64_t *counts) {
  return (int)guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    if (key_words < 1 || key_words > 2 || !hash_kind_supported(hash_kind) || p == 0 || p > PART_MAX_P || !counts) {
      set_error("bad partition arguments (1 <= p <= %d)", PART_MAX_P); return KH_ERR_INVALID;
    }
    if (n >= (1ULL << 32)) { set_error("partition batch must be < 2^32 keys"); return KH_ERR_INVALID; }
    DeviceGuard g(device);
    const int kind = fold_kind(hash_kind);
    const HashParams hp = make_hash_params(kind, seed, key_words);
    partition_dev(device, kind, hp, key_words, p, d_keys, n, d_out_keys, d_out_perm, counts, nullptr, 0);
    return KH_OK;
  });
}

int kh_partition_fast_dev(int device, int hash_kind, uint32_t seed, int key_words, uint32_t p,
                          const uint64_t *d_keys, uint64_t n, uint64_t *d_out_keys, uint32_t *d_out_perm,
                          uint64_t *counts) {
  return (int)guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    if (key_words < 1 || key_words > 2 || !hash_kind_supported(hash_kind) || p == 0 || p > (uint32_t)WIN_MAX_P || !counts) {
      set_error("bad partition arguments (1 <= p <= %d)", WIN_MAX_P); return KH_ERR_INVALID;
    }
    if (n >= (1ULL << 32)) { set_error("partition batch must be < 2^32 keys"); return KH_ERR_INVALID; }
    for (uint32_t r = 0; r < p; ++r) counts[r] = 0;
    if (n == 0) return KH_OK;
    DeviceGuard g(device);
    const int kind = fold_kind(hash_kind);
    const BinSpec spec{1, 0, p, kind, make_hash_params(kind, seed, key_words)};
    const int sms = sm_count(device);
    cudaStream_t s = cudaStreamPerThread;
    DevBuf<unsigned long long> hist(p), cursors(p), totals(p);
    KH_CUDA(cudaMemsetAsync(hist.p, 0, p * sizeof(unsigned long long), s));
    const int hgrid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 2047) / 2048, (uint64_t)sms * 4));
    if (key_words == 1) KH_LAUNCH((bin_hist_kernel<1>), hgrid, 512, 0, s, d_keys, n, spec, (int)p, hist.p);
    else KH_LAUNCH((bin_hist_kernel<2>), hgrid, 512, 0, s, d_keys, n, spec, (int)p, hist.p);
    KH_LAUNCH(bin_cursor_kernel, 1, REGION_FINE, 0, s, hist.p, (int)p, cursors.p, totals.p);
    std::vector<unsigned long long> h(p);
    KH_CUDA(cudaMemcpyAsync(h.data(), totals.p, p * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    if (key_words == 1) {
      if (d_out_perm) KH_LAUNCH((window_scatter_kernel<1, true>), sms * 4, WIN_BLOCK, 0, s, d_keys, n, spec, cursors.p, d_out_keys, d_out_perm);
      else KH_LAUNCH((window_scatter_kernel<1, false>), sms * 4, WIN_BLOCK, 0, s, d_keys, n, spec, cursors.p, d_out_keys, d_out_perm);
    } else {
      if (d_out_perm) KH_LAUNCH((window_scatter_kernel<2, true>), sms * 4, WIN_BLOCK, 0, s, d_keys, n, spec, cursors.p, d_out_keys, d_out_perm);
      else KH_LAUNCH((window_scatter_kernel<2, false>), sms * 4, WIN_BLOCK, 0, s, d_keys, n, spec, cursors.p, d_out_keys, d_out_perm);
    }
    KH_CUDA(cudaStreamSynchronize(s));
    for (uint32_t r = 0; r < p; ++r) counts[r] = h[r];
    return KH_OK;
  });
}

// kh_partition_fast_dev for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc): the same two kernels, their scratch -- histogram, cursors and totals, 3 x p counters -- in a file-scope device array
// instead of pool allocations.  Every GPU that loads the library holds the KH_PARTITION_NOALLOC_MAX_P * 24 bytes until exit.  All
// calls on a device share the array: the histogram is cleared at the start of every call, and calls that overlap on the GPU (two
// host threads, i.e. two per-thread streams) race on it.  No counterpart in the reference.
int kh_partition_fast_noalloc_dev(int device, int hash_kind, uint32_t seed, int key_words, uint32_t p,
                                  const uint64_t *d_keys, uint64_t n, uint64_t *d_out_keys, uint32_t *d_out_perm,
                                  uint64_t *counts) {
  return (int)guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    if (key_words < 1 || key_words > 2 || !hash_kind_supported(hash_kind) || p == 0 || !counts || (n && (!d_keys || !d_out_keys))) {
      set_error("bad partition arguments"); return KH_ERR_INVALID;
    }
    if (p > (uint32_t)KH_PARTITION_NOALLOC_MAX_P) {
      set_error("kh_partition_fast_noalloc_dev: p = %u exceeds the static scratch (at most %d destinations)", p, KH_PARTITION_NOALLOC_MAX_P);
      return KH_ERR_INVALID;
    }
    if (n >= (1ULL << 32)) { set_error("partition batch must be < 2^32 keys"); return KH_ERR_INVALID; }
    for (uint32_t r = 0; r < p; ++r) counts[r] = 0;
    if (n == 0) return KH_OK;
    DeviceGuard g(device);
    unsigned long long *scratch = noalloc_scratch(device);          // device address of g_part_scratch on THIS device
    unsigned long long *hist = scratch, *cursors = scratch + KH_PARTITION_NOALLOC_MAX_P, *totals = scratch + 2 * KH_PARTITION_NOALLOC_MAX_P;
    const int kind = fold_kind(hash_kind);
    const BinSpec spec{1, 0, p, kind, make_hash_params(kind, seed, key_words)};
    const int sms = sm_count(device);
    cudaStream_t s = cudaStreamPerThread;
    KH_CUDA(cudaMemsetAsync(hist, 0, p * sizeof(unsigned long long), s));             // the array keeps the last call's contents
    const int hgrid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 2047) / 2048, (uint64_t)sms * 4));
    if (key_words == 1) KH_LAUNCH((bin_hist_kernel<1>), hgrid, 512, 0, s, d_keys, n, spec, (int)p, hist);
    else KH_LAUNCH((bin_hist_kernel<2>), hgrid, 512, 0, s, d_keys, n, spec, (int)p, hist);
    KH_LAUNCH(bin_cursor_kernel, 1, REGION_FINE, 0, s, hist, (int)p, cursors, totals);
    unsigned long long h[KH_PARTITION_NOALLOC_MAX_P];
    KH_CUDA(cudaMemcpyAsync(h, totals, p * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    if (key_words == 1) {
      if (d_out_perm) KH_LAUNCH((window_scatter_kernel<1, true>), sms * 4, WIN_BLOCK, 0, s, d_keys, n, spec, cursors, d_out_keys, d_out_perm);
      else KH_LAUNCH((window_scatter_kernel<1, false>), sms * 4, WIN_BLOCK, 0, s, d_keys, n, spec, cursors, d_out_keys, d_out_perm);
    } else {
      if (d_out_perm) KH_LAUNCH((window_scatter_kernel<2, true>), sms * 4, WIN_BLOCK, 0, s, d_keys, n, spec, cursors, d_out_keys, d_out_perm);
      else KH_LAUNCH((window_scatter_kernel<2, false>), sms * 4, WIN_BLOCK, 0, s, d_keys, n, spec, cursors, d_out_keys, d_out_perm);
    }
    KH_CUDA(cudaStreamSynchronize(s));
    for (uint32_t r = 0; r < p; ++r) counts[r] = h[r];
    return KH_OK;
  });
}

int kh_rank_host(int device, int hash_kind, uint32_t seed, int key_words, uint32_t p,
                 const uint64_t *keys, uint64_t n, uint32_t *out_rank) {
  return (int)guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    if (key_words < 1 || key_words > 2 || !hash_kind_supported(hash_kind) || p == 0) { set_error("bad rank arguments"); return KH_ERR_INVALID; }
    if (n == 0) return KH_OK;
    DeviceGuard g(device);
    const int kind = fold_kind(hash_kind);
    const HashParams hp = make_hash_params(kind, seed, key_words);
    DevBuf<uint64_t> dk(n * key_words); DevBuf<uint32_t> dr(n);
    KH_CUDA(cudaMemcpy(dk.p, keys, n * key_words * 8, cudaMemcpyHostToDevice));
    const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count(device) * 8));
    if (key_words == 1) KH_LAUNCH((rank_kernel<1>), grid, 256, 0, 0, dk.p, n, kind, hp, p, dr.p);
    else KH_LAUNCH((rank_kernel<2>), grid, 256, 0, 0, dk.p, n, kind, hp, p, dr.p);
    KH_CUDA(cudaMemcpy(out_rank, dr.p, n * 4, cudaMemcpyDeviceToHost));
    return KH_OK;
  });
}

int kh_unpermute_dev(int device, int elem_bytes, const void *d_in, const uint32_t *d_perm, uint64_t n, void *d_out) {
  return (int)guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    if (elem_bytes != 1 && elem_bytes != 4) { set_error("elem_bytes must be 1 or 4"); return KH_ERR_INVALID; }
    if (n == 0) return KH_OK;
    DeviceGuard g(device);
    const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count(device) * 8));
    if (elem_bytes == 1) KH_LAUNCH((unpermute_kernel<uint8_t>), grid, 256, 0, 0, (const uint8_t *)d_in, d_perm, n, (uint8_t *)d_out);
    else KH_LAUNCH((unpermute_kernel<uint32_t>), grid, 256, 0, 0, (const uint32_t *)d_in, d_perm, n, (uint32_t *)d_out);
    KH_CUDA(cudaStreamSynchronize(0));
    return KH_OK;
  });
}

}  // extern "C"
