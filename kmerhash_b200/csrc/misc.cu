// misc.cu -- error state, launch accounting, hash parameter setup, pinned memory, hash / HLL C ABI.
#include <cstdarg>
#include <cstring>
#include <mutex>
#include "kh_common.h"
#include "kh_hash.cuh"
#include "kh_table.cuh"

namespace kh {

static thread_local char g_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char *fmt, ...) {
  va_list ap; va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count(int device) {
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
  return n;
}

void pool_setup(int device) {
  static std::mutex mu;
  static bool done[64] = {false};
  if (device < 0 || device >= 64) return;
  std::lock_guard<std::mutex> lk(mu);
  if (done[device]) return;
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
    uint64_t keep = ~0ULL;   // never trim: freed scratch stays with the pool for the next batch
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  cudaGetLastError();
  done[device] = true;
}

bool hash_kind_supported(int kind) {
  // murmur3finalizer_avx32 static_asserts sizeof(T) <= 4 (KH/murmurhash3finalizer_32_avx.hpp:527):
  // it cannot be instantiated for 8/16-byte k-mers in the reference either.
  return kind >= KH_HASH_IDEN && kind <= KH_HASH_MURMUR_X86 && kind != KH_HASH_MURMUR32FIN;
}

// glibc rand() after srand(seed): the TYPE_3 additive-feedback generator (degree 31, separation 3),
// state filled by the Park-Miller LCG and warmed up 310 steps.  fsc::hash::clhash derives its 133
// key words from it as (rand() << 32) | rand()   (KH/hash_new.hpp:288-296).
namespace {
struct GlibcRand {
  uint32_t ring[31];
  int front, rear;
  explicit GlibcRand(uint32_t seed) {
    int64_t x = seed ? seed : 1;
    ring[0] = (uint32_t)x;
    for (int i = 1; i < 31; ++i) {
      x = (16807 * x) % 2147483647;          // Park-Miller; x stays positive in 64-bit arithmetic
      ring[i] = (uint32_t)x;
    }
    front = 3; rear = 0;
    for (int i = 0; i < 310; ++i) next_raw();
  }
  uint32_t next_raw() {
    ring[front] += ring[rear];
    uint32_t r = ring[front];
    front = (front + 1) % 31; rear = (rear + 1) % 31;
    return r;
  }
  uint32_t next() { return next_raw() >> 1; }
};
}  // namespace

HashParams make_hash_params(int kind, uint32_t seed, int key_words) {
  HashParams p;
  p.seed = seed; p.cl0 = p.cl1 = p.cllen_lo = p.cllen_hi = 0;
  if (fold_kind(kind) == KH_HASH_CLHASH) {
    GlibcRand g(seed);
    uint64_t k0 = 0, k1 = 0, k132 = 0;
    for (int i = 0; i < 133; ++i) {
      const uint64_t hi = g.next(), lo = g.next();
      const uint64_t w = (hi << 32) | lo;
      if (i == 0) k0 = w; else if (i == 1) k1 = w; else if (i == 132) k132 = w;
    }
    p.cl0 = k0; p.cl1 = k1;
    clmul64((uint64_t)(8 * key_words), k132, p.cllen_lo, p.cllen_hi);   // lazyLengthHash, clhash.c:81-85
  }
  return p;
}

template <int W>
__global__ void hash_batch_kernel(const uint64_t *keys, uint64_t n, int kind, HashParams hp, uint64_t *out) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t key[W];
    load_key<W>(keys, i, key);
    out[i] = hash_runtime<W>(kind, key, hp);
  }
}

}  // namespace kh

using namespace kh;

extern "C" {

const char *kh_last_error(void) { return g_err; }
int kh_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
uint64_t kh_launch_count(void) { return g_launches.load(); }
void kh_reset_launch_count(void) { g_launches.store(0); }

void *kh_host_alloc(size_t bytes) {
  void *p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
    set_error("cudaHostAlloc(%zu) failed", bytes); cudaGetLastError(); return nullptr;
  }
  return p;
}
void kh_host_free(void *p) { if (p) cudaFreeHost(p); }

int kh_hash_dev(int device, int hash_kind, uint32_t seed, int key_words, const uint64_t *d_keys, uint64_t n, uint64_t *d_out) {
  return (int)guarded([&]() -> int64_t {
    if (key_words < 1 || key_words > 2 || !hash_kind_supported(hash_kind)) { set_error("unsupported hash kind / key width"); return KH_ERR_INVALID; }
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    if (n == 0) return KH_OK;
    DeviceGuard g(device);
    const int kind = fold_kind(hash_kind);
    const HashParams hp = make_hash_params(kind, seed, key_words);
    const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((n + 255) / 256, (uint64_t)sm_count(device) * 8));
    if (key_words == 1) KH_LAUNCH((hash_batch_kernel<1>), grid, 256, 0, 0, d_keys, n, kind, hp, d_out);
    else KH_LAUNCH((hash_batch_kernel<2>), grid, 256, 0, 0, d_keys, n, kind, hp, d_out);
    KH_CUDA(cudaStreamSynchronize(0));
    return KH_OK;
  });
}

int kh_hash_host(int device, int hash_kind, uint32_t seed, int key_words, const uint64_t *keys, uint64_t n, uint64_t *out) {
  return (int)guarded([&]() -> int64_t {
    if (key_words < 1 || key_words > 2 || !hash_kind_supported(hash_kind)) { set_error("unsupported hash kind / key width"); return KH_ERR_INVALID; }
    if (kh_device_count() <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    if (n == 0) return KH_OK;
    DeviceGuard g(device);
    DevBuf<uint64_t> dk(n * key_words), dout(n);
    KH_CUDA(cudaMemcpy(dk.p, keys, n * key_words * 8, cudaMemcpyHostToDevice));
    const int rc = kh_hash_dev(device, hash_kind, seed, key_words, dk.p, n, dout.p);
    if (rc != KH_OK) return rc;
    KH_CUDA(cudaMemcpy(out, dout.p, n * 8, cudaMemcpyDeviceToHost));
    return KH_OK;
  });
}

}  // extern "C"
