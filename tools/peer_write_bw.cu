// Peer-store bandwidth of one GPU into another over NVLink, by access shape (what bounds the multi-GPU sender):
//   contig16 / contig8   contiguous, 16 / 8 bytes per lane
//   runsRx8B             runs of R keys (8 B) that start 8 bytes off a 16-byte boundary, 8 bytes per lane: every warp store
//                        straddles two 128-byte lines (the sender's copy-out before the runs were line-aligned)
//   runsRx16B            the same runs, 16 bytes per lane behind a scalar head
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o peer_write_bw peer_write_bw.cu ; needs 2 GPUs
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void copy16(const uint4 *src, uint4 *dst, size_t n16) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
__global__ void copy8(const uint64_t *src, uint64_t *dst, size_t n8) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
// run r goes to dst + r * (R + 1) + 1: an odd start, never 16-byte aligned twice in a row
__global__ void runs8(const uint64_t *src, uint64_t *dst, size_t n_runs, int R) {
  for (size_t r = blockIdx.x; r < n_runs; r += gridDim.x) {
    const uint64_t *s = src + r * R;
    uint64_t *d = dst + r * (size_t)(R + 1) + 1;
    for (int i = threadIdx.x; i < R; i += blockDim.x) d[i] = s[i];
  }
}
__global__ void runs16(const uint64_t *src, uint64_t *dst, size_t n_runs, int R) {
  for (size_t r = blockIdx.x; r < n_runs; r += gridDim.x) {
    const uint64_t *s = src + r * R;
    uint64_t *d = dst + r * (size_t)(R + 1) + 1;
    const int head = (int)(((uintptr_t)d >> 3) & 1);
    if (threadIdx.x == 0 && head) d[0] = s[0];
    const int pairs = (R - head) / 2;
    for (int i = threadIdx.x; i < pairs; i += blockDim.x)
      *reinterpret_cast<ulonglong2 *>(d + head + 2 * i) = make_ulonglong2(s[head + 2 * i], s[head + 2 * i + 1]);
    if (threadIdx.x == 0 && head + 2 * pairs < R) d[R - 1] = s[R - 1];
  }
}

int main() {
  int nd = 0; CK(cudaGetDeviceCount(&nd));
  if (nd < 2) { printf("needs 2 GPUs\n"); return 0; }
  const size_t bytes = 800ull << 20;
  uint64_t *src, *loc, *peer;
  CK(cudaSetDevice(1)); CK(cudaMalloc(&peer, bytes + (64 << 20)));
  CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0));
  CK(cudaMalloc(&src, bytes)); CK(cudaMalloc(&loc, bytes + (64 << 20))); CK(cudaMemset(src, 1, bytes));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int grids[3] = {148, 148 * 3, 148 * 8};
  const char *names[8] = {"contig16", "contig8", "runs256x8B", "runs256x16B", "runs1024x8B", "runs1024x16B", "runs4096x8B", "runs4096x16B"};
  for (int tgt = 0; tgt < 2; ++tgt) {
    uint64_t *dst = tgt ? peer : loc;
    for (int gi = 0; gi < 3; ++gi) {
      const int grid = grids[gi];
      for (int mode = 0; mode < 8; ++mode) {
        const int R = mode == 2 || mode == 3 ? 256 : (mode == 4 || mode == 5 ? 1024 : 4096);
        const size_t n_runs = bytes / 8 / R;
        float best = 1e9f;
        for (int it = 0; it < 4; ++it) {
          CK(cudaEventRecord(e0));
          if (mode == 0) copy16<<<grid, 512>>>((const uint4 *)src, (uint4 *)dst, bytes / 16);
          else if (mode == 1) copy8<<<grid, 512>>>(src, dst, bytes / 8);
          else if (mode & 1) runs16<<<grid, 512>>>(src, dst, n_runs, R);
          else runs8<<<grid, 512>>>(src, dst, n_runs, R);
          CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
          float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
          if (it && ms < best) best = ms;
        }
        printf("{\"target\": \"%s\", \"grid\": %d, \"shape\": \"%s\", \"ms\": %.4f, \"GBps\": %.1f}\n", tgt ? "peer" : "local", grid, names[mode], best,
               bytes / best / 1e6);
      }
    }
  }
  return 0;
}
