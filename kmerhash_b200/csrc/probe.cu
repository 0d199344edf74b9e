// probe.cu -- roofline probes (measurement support; SURVEY.md 8d "measured random 32-B-sector roofline").
// kh_probe_random_access touches uniformly random, aligned pieces of a table-sized buffer, either
// read-only (find/count ceiling) or read + atomicAdd on the first word (insert/erase ceiling).
// kh_probe_stream_copy is the plain copy the HBM peak in MEASURED_PEAKS.json was taken with.
#include <algorithm>
#include "kh_common.h"

namespace kh {

__device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z += 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}

template <int BYTES, int MODE>
__global__ void __launch_bounds__(256) random_access_kernel(unsigned char *buf, uint64_t n_pieces_mask, uint64_t n_access,
                                                            uint64_t salt, unsigned long long *sink) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t acc = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_access; i += stride) {
    const uint64_t piece = mix64(i ^ salt) & n_pieces_mask;
    unsigned char *p = buf + piece * BYTES;
    if (BYTES == 16) {
      uint64_t a, b;
      asm volatile("ld.global.cg.v2.u64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p) : "memory");
      acc += a ^ b;
    } else {
#pragma unroll
      for (int o = 0; o < BYTES; o += 32) {
        uint64_t a, b, c, d;
        asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p + o) : "memory");
        acc += a ^ b ^ c ^ d;
      }
    }
    if (MODE == 1) atomicAdd(reinterpret_cast<unsigned int *>(p + 8), 1u);
  }
  if (acc == 0x123456789ULL) atomicAdd(sink, 1ULL);   // keep the loads alive
}

// same traffic, four independent accesses in flight per thread (does more memory-level parallelism
// per thread raise the ceiling?)
template <int BYTES, int MODE>
__global__ void __launch_bounds__(256) random_access_ilp4_kernel(unsigned char *buf, uint64_t n_pieces_mask, uint64_t n_access,
                                                                 uint64_t salt, unsigned long long *sink) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  uint64_t acc = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_access; i += 4 * stride) {
    unsigned char *p[4];
    uint64_t w[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u) p[u] = buf + (mix64((i + u * stride) ^ salt) & n_pieces_mask) * BYTES;
#pragma unroll
    for (int o = 0; o < BYTES; o += 32) {
#pragma unroll
      for (int u = 0; u < 4; ++u)
        asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(w[u][0]), "=l"(w[u][1]), "=l"(w[u][2]), "=l"(w[u][3]) : "l"(p[u] + o));
#pragma unroll
      for (int u = 0; u < 4; ++u) acc += w[u][0] ^ w[u][1] ^ w[u][2] ^ w[u][3];
    }
    if (MODE == 1) {
#pragma unroll
      for (int u = 0; u < 4; ++u) if (i + u * stride < n_access) atomicAdd(reinterpret_cast<unsigned int *>(p[u] + 8), 1u);
    }
  }
  if (acc == 0x123456789ULL) atomicAdd(sink, 1ULL);
}

__global__ void stream_copy_kernel(const uint4 *__restrict__ in, uint4 *__restrict__ out, uint64_t n) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = in[i];
}

}  // namespace kh

using namespace kh;

extern "C" {

double kh_probe_random_access(int device, uint64_t buf_bytes, uint64_t n_access, int bytes_per_access, int mode, int iters) {
  double ms_out = -1.0;
  guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device"); return KH_ERR_CUDA; }
    const bool ilp4 = mode >= 2;   // mode 2 = read, 3 = read + atomicAdd, four accesses in flight per thread
    if (ilp4) mode -= 2;
    if (ilp4 && bytes_per_access == 16) { set_error("ilp4 probe: 32, 64 or 128 bytes"); return KH_ERR_INVALID; }
    if (bytes_per_access != 16 && bytes_per_access != 32 && bytes_per_access != 64 && bytes_per_access != 128) {
      set_error("bytes_per_access must be 16, 32, 64 or 128"); return KH_ERR_INVALID;
    }
    DeviceGuard g(device);
    uint64_t pieces = 1; while (pieces * 2 * bytes_per_access <= buf_bytes) pieces *= 2;
    DevBuf<unsigned char> buf(pieces * bytes_per_access);
    DevBuf<unsigned long long> sink(1);
    KH_CUDA(cudaMemset(buf.p, 0, pieces * bytes_per_access));
    KH_CUDA(cudaMemset(sink.p, 0, 8));
    const int grid = sm_count(device) * 8;
    cudaEvent_t e0, e1;
    KH_CUDA(cudaEventCreate(&e0)); KH_CUDA(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int it = 0; it < std::max(iters, 1) + 1; ++it) {
      const uint64_t salt = 0x5bd1e995ULL * (it + 1);
      KH_CUDA(cudaEventRecord(e0, 0));
#define KH_RA4(B)                                                                                             \
      if (mode == 0) KH_LAUNCH((random_access_ilp4_kernel<B, 0>), grid, 256, 0, 0, buf.p, pieces - 1, n_access, salt, sink.p); \
      else KH_LAUNCH((random_access_ilp4_kernel<B, 1>), grid, 256, 0, 0, buf.p, pieces - 1, n_access, salt, sink.p);
      if (ilp4) { switch (bytes_per_access) { case 32: KH_RA4(32) break; case 64: KH_RA4(64) break; default: KH_RA4(128) break; } }
      else
#define KH_RA(B)                                                                                              \
      if (mode == 0) KH_LAUNCH((random_access_kernel<B, 0>), grid, 256, 0, 0, buf.p, pieces - 1, n_access, salt, sink.p); \
      else KH_LAUNCH((random_access_kernel<B, 1>), grid, 256, 0, 0, buf.p, pieces - 1, n_access, salt, sink.p);
      switch (bytes_per_access) { case 16: KH_RA(16) break; case 32: KH_RA(32) break; case 64: KH_RA(64) break; default: KH_RA(128) break; }
#undef KH_RA
#undef KH_RA4
      KH_CUDA(cudaEventRecord(e1, 0));
      KH_CUDA(cudaEventSynchronize(e1));
      float ms = 0; KH_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      if (it > 0) best = std::min(best, ms);   // first iteration is warm-up
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    ms_out = best;
    return KH_OK;
  });
  return ms_out;
}

double kh_probe_stream_copy(int device, uint64_t bytes, int iters) {
  double ms_out = -1.0;
  guarded([&]() -> int64_t {
    if (kh_device_count() <= 0) { set_error("no CUDA device"); return KH_ERR_CUDA; }
    DeviceGuard g(device);
    const uint64_t n = bytes / 16;
    DevBuf<uint4> a(n), b(n);
    KH_CUDA(cudaMemset(a.p, 1, n * 16));
    cudaEvent_t e0, e1;
    KH_CUDA(cudaEventCreate(&e0)); KH_CUDA(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int it = 0; it < std::max(iters, 1) + 1; ++it) {
      KH_CUDA(cudaEventRecord(e0, 0));
      KH_LAUNCH(stream_copy_kernel, sm_count(device) * 8, 256, 0, 0, a.p, b.p, n);
      KH_CUDA(cudaEventRecord(e1, 0));
      KH_CUDA(cudaEventSynchronize(e1));
      float ms = 0; KH_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      if (it > 0) best = std::min(best, ms);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    ms_out = best;
    return KH_OK;
  });
  return ms_out;
}

}  // extern "C"
