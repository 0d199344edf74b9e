// kh_common.h -- host-side plumbing shared by the translation units of libkmerhash_b200.so:
// error state, CUDA checks, launch accounting, device buffers.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <string>
#include <atomic>
#include "../../include/kmerhash_b200.h"

namespace kh {

void set_error(const char *fmt, ...);
extern std::atomic<uint64_t> g_launches;

struct CudaError { int code; };

#define KH_CUDA(expr)                                                                         \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      kh::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, cudaGetErrorString(_e)); \
      throw kh::CudaError{_e == cudaErrorMemoryAllocation ? KH_ERR_NOMEM : KH_ERR_CUDA};      \
    }                                                                                         \
  } while (0)

// every kernel launch of the library goes through this macro so kh_launch_count() is honest
#define KH_LAUNCH(kernel, grid, block, smem, stream, ...)                                     \
  do {                                                                                        \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                               \
    kh::g_launches.fetch_add(1, std::memory_order_relaxed);                                   \
    KH_CUDA(cudaGetLastError());                                                              \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    KH_CUDA(cudaGetDevice(&prev));
    if (prev != dev) KH_CUDA(cudaSetDevice(dev));
    else prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// RAII device allocation from the device's stream-ordered memory pool (cudaMallocAsync).  The pool keeps
// freed blocks (release threshold = max, set once per device in pool_setup), so the scratch arrays of
// every batch call (hundreds of MB: sorted keys, staging, rebuild dumps) cost microseconds instead of
// the milliseconds cudaMalloc/cudaFree take.  alloc() returns memory that is safe on ANY stream (the
// allocating stream is synchronised); release() requires that the work using the buffer has completed,
// which every C-ABI entry point guarantees by synchronising its streams before its buffers go out of scope.
void pool_setup(int device);
template <typename T>
struct DevBuf {
  T *p = nullptr;
  size_t n = 0;
  DevBuf() {}
  explicit DevBuf(size_t count) { alloc(count); }
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  DevBuf(DevBuf &&o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
  DevBuf &operator=(DevBuf &&o) noexcept { if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; } return *this; }
  void alloc(size_t count) {
    release();
    n = count;
    if (count) {
      int dev = 0;
      KH_CUDA(cudaGetDevice(&dev));
      pool_setup(dev);
      KH_CUDA(cudaMallocAsync((void **)&p, count * sizeof(T), cudaStreamPerThread));
      KH_CUDA(cudaStreamSynchronize(cudaStreamPerThread));
    }
  }
  void ensure(size_t count) { if (count > n) alloc(count); }
  void release() { if (p) { cudaFreeAsync(p, cudaStreamPerThread); p = nullptr; n = 0; } }
  ~DevBuf() { release(); }
};

int sm_count(int device);
inline uint64_t next_pow2(uint64_t x) { uint64_t p = 1; while (p < x) p <<= 1; return p; }

// C-ABI wrapper: converts exceptions into negative status codes
template <typename F>
inline int64_t guarded(F &&f) {
  try { return f(); }
  catch (const CudaError &e) { return e.code; }
  catch (const std::bad_alloc &) { set_error("host allocation failed"); return KH_ERR_NOMEM; }
  catch (const std::exception &e) { set_error("%s", e.what()); return KH_ERR_INVALID; }
}

}  // namespace kh
