// dist.cu -- khd_map: one table per GPU of the box, driven by one host thread group.
//
// Replaces the reference's distributed containers and their MPI exchange:
//   dsc::batched_robinhood_map_base::insert_p / count_p / find_p / erase_p
//       KH/distributed_batched_robinhood_map.hpp:910-1179,1258-1516,1619,2169
//   counting_batched_robinhood_map (inserts T(1) under std::plus)        :2542-2649
//   transform_input (lex_less canonicalisation)   KI/containers/distributed_map_base.hpp:287-301
//   mxx::all2all(counts) + khmxx::distribute_permuted -> MPI_Alltoallv   KH/incremental_mxx.hpp:2998-3031
// Owner of a key = DistHash_{9876543}(canonical key) & (p-1) (or % p), the reference's own rule,
// so every GPU's table holds exactly what the reference's rank of the same index holds.
//
// B200 shape: the input batch is split evenly over the GPUs (the "local input" of each rank); each
// GPU histograms its slice, the p x p count matrix is combined on the host (the all2all of counts is
// p*p integers), and then the stable scatter kernel of GPU g stores every key DIRECTLY into the
// receive buffer of its owner GPU through peer-mapped pointers over NVLink -- partition and
// all-to-all are one kernel, there is no send staging buffer and no separate exchange step.
// Queries send their 1-byte / 4-byte answers back with peer copies and are un-permuted on the
// source GPU, so results come back in INPUT order.  Without peer access (no NVLink) the same
// schedule runs with cudaMemcpyPeerAsync per segment.
#include <algorithm>
#include <functional>
#include <thread>
#include <vector>
#include "kh_common.h"
#include "kh_partition.h"

using namespace kh;

struct khd_map {
  int p = 1, W = 1, store_hash = 0, dist_hash = 0, reducer = 0, canon_k = 0;
  std::vector<int> dev;
  std::vector<kh_table *> table;
  std::vector<cudaStream_t> stream;
  HashParams dist_hp;
  bool p2p = false;
};

namespace {

void parallel_for(int n, const std::function<void(int)> &fn) {
  if (n == 1) { fn(0); return; }
  std::vector<std::thread> th;
  std::vector<int> err(n, 0);
  std::vector<std::string> msg(n);
  for (int i = 0; i < n; ++i) th.emplace_back([&, i]() {
    try { fn(i); }
    catch (const CudaError &e) { err[i] = e.code; msg[i] = kh_last_error(); }
    catch (const std::exception &e) { err[i] = KH_ERR_INVALID; msg[i] = e.what(); }
  });
  for (auto &t : th) t.join();
  for (int i = 0; i < n; ++i) if (err[i]) { set_error("gpu %d: %s", i, msg[i].c_str()); throw CudaError{err[i]}; }
}

// Everything one batch operation needs after the keys have reached their owners.
struct Exchange {
  int p = 1, W = 1;
  std::vector<uint64_t> slice_beg;                 // [p+1] input slice of every source GPU
  std::vector<std::vector<uint64_t>> C;            // C[g][r] keys GPU g sends to GPU r
  std::vector<std::vector<uint64_t>> recv_off;     // recv_off[r][g] offset of g's segment in r's receive buffer
  std::vector<uint64_t> recv_n;                    // keys received by r
  std::vector<DevBuf<uint64_t>> d_in, d_recv;      // per GPU: its input slice, its received keys
  std::vector<DevBuf<uint32_t>> d_perm;            // per source GPU: original index of every permuted position
  std::vector<PartitionPlan> plan;
};

void run_exchange(khd_map *m, const uint64_t *keys, uint64_t n, bool want_perm, Exchange &x) {
  const int p = m->p, W = m->W;
  x.p = p; x.W = W;
  x.slice_beg.resize(p + 1);
  for (int g = 0; g <= p; ++g) x.slice_beg[g] = n / p * g + std::min<uint64_t>(g, n % p);
  x.C.assign(p, std::vector<uint64_t>(p, 0));
  x.recv_off.assign(p, std::vector<uint64_t>(p, 0));
  x.recv_n.assign(p, 0);
  x.d_in.resize(p); x.d_recv.resize(p); x.d_perm.resize(p); x.plan.resize(p);

  // 1. slices to their GPUs, canonicalise, histogram by owner
  parallel_for(p, [&](int g) {
    DeviceGuard dg(m->dev[g]);
    cudaStream_t s = m->stream[g];
    const uint64_t ng = x.slice_beg[g + 1] - x.slice_beg[g];
    x.d_in[g].alloc(std::max<uint64_t>(ng * W, 1));
    if (ng) KH_CUDA(cudaMemcpyAsync(x.d_in[g].p, keys + x.slice_beg[g] * W, ng * W * 8, cudaMemcpyHostToDevice, s));
    KH_CUDA(cudaStreamSynchronize(s));
    if (m->canon_k > 0 && ng) {
      const int rc = kh_canonicalize_dev(m->dev[g], m->canon_k, x.d_in[g].p, ng, x.d_in[g].p);
      if (rc != KH_OK) throw CudaError{rc};
    }
    if (p == 1) { x.C[0][0] = ng; return; }
    x.plan[g].init(m->dev[g], m->dist_hash, m->dist_hp, W, (uint32_t)p, x.d_in[g].p, ng);
    x.plan[g].count(s, x.C[g].data());
  });
  // 2. the "all2all of counts": combine the p x p matrix on the host
  for (int r = 0; r < p; ++r) {
    uint64_t o = 0;
    for (int g = 0; g < p; ++g) { x.recv_off[r][g] = o; o += x.C[g][r]; }
    x.recv_n[r] = o;
  }
  if (p == 1) {          // insert_1 / count_1 path of the reference: no partition, no exchange
    x.d_recv[0].p = x.d_in[0].p; x.d_recv[0].n = x.d_in[0].n; x.d_in[0].p = nullptr; x.d_in[0].n = 0;
    return;
  }
  parallel_for(p, [&](int r) {
    DeviceGuard dg(m->dev[r]);
    x.d_recv[r].alloc(std::max<uint64_t>(x.recv_n[r] * W, 1));
  });
  // 3. scatter = partition + exchange in one kernel (peer stores), or scatter + peer copies
  parallel_for(p, [&](int g) {
    DeviceGuard dg(m->dev[g]);
    cudaStream_t s = m->stream[g];
    const uint64_t ng = x.slice_beg[g + 1] - x.slice_beg[g];
    if (want_perm) x.d_perm[g].alloc(std::max<uint64_t>(ng, 1));
    if (ng == 0) return;
    if (m->p2p) {
      PartDst dst{};
      for (int r = 0; r < p; ++r) dst.base[r] = x.d_recv[r].p + x.recv_off[r][g] * W;
      x.plan[g].scatter(s, nullptr, want_perm ? x.d_perm[g].p : nullptr, &dst);
    } else {
      DevBuf<uint64_t> tmp(ng * W);
      x.plan[g].scatter(s, tmp.p, want_perm ? x.d_perm[g].p : nullptr, nullptr);
      for (int r = 0; r < p; ++r) if (x.C[g][r])
        KH_CUDA(cudaMemcpyPeerAsync(x.d_recv[r].p + x.recv_off[r][g] * W, m->dev[r], tmp.p + x.plan[g].seg_h[r] * W,
                                    m->dev[g], x.C[g][r] * W * 8, s));
      KH_CUDA(cudaStreamSynchronize(s));
    }
    KH_CUDA(cudaStreamSynchronize(s));
  });
}

// answers of GPU r's local lookup travel back to the source GPUs and into input order
template <typename T>
void return_results(khd_map *m, Exchange &x, std::vector<DevBuf<T>> &d_res, T *out) {
  const int p = m->p;
  parallel_for(p, [&](int g) {
    DeviceGuard dg(m->dev[g]);
    cudaStream_t s = m->stream[g];
    const uint64_t ng = x.slice_beg[g + 1] - x.slice_beg[g];
    if (ng == 0) return;
    if (p == 1) {
      KH_CUDA(cudaMemcpyAsync(out, d_res[0].p, ng * sizeof(T), cudaMemcpyDeviceToHost, s));
      KH_CUDA(cudaStreamSynchronize(s));
      return;
    }
    DevBuf<T> perm_order(ng), in_order(ng);
    for (int r = 0; r < p; ++r) if (x.C[g][r])
      KH_CUDA(cudaMemcpyPeerAsync(perm_order.p + x.plan[g].seg_h[r], m->dev[g], d_res[r].p + x.recv_off[r][g], m->dev[r],
                                  x.C[g][r] * sizeof(T), s));
    KH_CUDA(cudaStreamSynchronize(s));
    const int rc = kh_unpermute_dev(m->dev[g], (int)sizeof(T), perm_order.p, x.d_perm[g].p, ng, in_order.p);
    if (rc != KH_OK) throw CudaError{rc};
    KH_CUDA(cudaMemcpy(out + x.slice_beg[g], in_order.p, ng * sizeof(T), cudaMemcpyDeviceToHost));
  });
}

// values follow their keys: pv[i] = v[perm[i]] (perm = original index of every permuted position)
__global__ void gather_vals_kernel(const uint32_t *v, const uint32_t *perm, uint64_t n, uint32_t *pv) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) pv[i] = v[perm[i]];
}

}  // namespace

extern "C" {

khd_map *khd_create(int n_gpus, int key_words, int store_hash, int dist_hash, int reducer, int canonical_k) {
  khd_map *m = nullptr;
  const int64_t rc = guarded([&]() -> int64_t {
    const int ndev = kh_device_count();
    if (ndev <= 0) { set_error("no CUDA device: kmerhash_b200 has no CPU fallback"); return KH_ERR_CUDA; }
    if (n_gpus < 1 || n_gpus > 8 || n_gpus > ndev) { set_error("n_gpus must be in [1, min(8, %d)]", ndev); return KH_ERR_INVALID; }
    if (key_words < 1 || key_words > 2 || !hash_kind_supported(store_hash) || !hash_kind_supported(dist_hash)) {
      set_error("bad key width / hash kind"); return KH_ERR_INVALID;
    }
    if (canonical_k < 0 || canonical_k > 32 * key_words) { set_error("canonical_k does not fit the key width"); return KH_ERR_INVALID; }
    m = new khd_map();
    m->p = n_gpus; m->W = key_words; m->store_hash = store_hash; m->dist_hash = fold_kind(dist_hash);
    m->reducer = reducer; m->canon_k = canonical_k;
    m->dist_hp = make_hash_params(m->dist_hash, KH_DIST_SEED, key_words);
    m->dev.resize(n_gpus); m->table.assign(n_gpus, nullptr); m->stream.assign(n_gpus, nullptr);
    m->p2p = n_gpus > 1;
    for (int g = 0; g < n_gpus; ++g) {
      m->dev[g] = g;
      DeviceGuard dg(g);
      KH_CUDA(cudaStreamCreateWithFlags(&m->stream[g], cudaStreamNonBlocking));
      for (int r = 0; r < n_gpus; ++r) if (r != g) {
        int can = 0;
        KH_CUDA(cudaDeviceCanAccessPeer(&can, g, r));
        if (!can) { m->p2p = false; continue; }
        cudaError_t e = cudaDeviceEnablePeerAccess(r, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) { cudaGetLastError(); m->p2p = false; }
      }
      // the receive buffers come from this device's stream-ordered pool (DevBuf): cudaDeviceEnablePeerAccess covers
      // cudaMalloc memory only, pool memory needs its own grant for every peer that stores into it
      if (n_gpus > 1) {
        pool_setup(g);
        cudaMemPool_t pool;
        KH_CUDA(cudaDeviceGetDefaultMemPool(&pool, g));
        std::vector<cudaMemAccessDesc> acc;
        for (int r = 0; r < n_gpus; ++r) if (r != g) {
          int can = 0;
          KH_CUDA(cudaDeviceCanAccessPeer(&can, r, g));
          if (!can) continue;
          cudaMemAccessDesc d{};
          d.location.type = cudaMemLocationTypeDevice; d.location.id = r; d.flags = cudaMemAccessFlagsProtReadWrite;
          acc.push_back(d);
        }
        if (!acc.empty() && cudaMemPoolSetAccess(pool, acc.data(), acc.size()) != cudaSuccess) { cudaGetLastError(); m->p2p = false; }
      }
      const uint32_t seed = (fold_kind(store_hash) == KH_HASH_CRC32C) ? KH_DEFAULT_CRC_SEED : KH_DEFAULT_SEED;
      m->table[g] = kh_create(g, KH_MAP_ROBINHOOD_OFFSETS, key_words, store_hash, seed, reducer);
      if (!m->table[g]) return KH_ERR_CUDA;
    }
    return KH_OK;
  });
  if (rc != KH_OK) { if (m) khd_destroy(m); return nullptr; }
  return m;
}

void khd_destroy(khd_map *m) {
  if (!m) return;
  int prev = -1;
  cudaGetDevice(&prev);                 // the caller's current device is not ours to change
  for (int g = 0; g < (int)m->table.size(); ++g) {
    if (m->table[g]) kh_destroy(m->table[g]);
    if (m->stream[g]) { cudaSetDevice(m->dev[g]); cudaStreamDestroy(m->stream[g]); }
  }
  if (prev >= 0) cudaSetDevice(prev);
  delete m;
}

int khd_n_gpus(const khd_map *m) { return m ? m->p : 0; }
int khd_uses_peer_stores(const khd_map *m) { return (m && m->p2p) ? 1 : 0; }

int64_t khd_insert_host(khd_map *m, const uint64_t *keys, uint64_t n, uint32_t default_val) {
  return guarded([&]() -> int64_t {
    if (!m || (n && !keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    Exchange x;
    run_exchange(m, keys, n, false, x);
    std::vector<int64_t> added(m->p, 0);
    // every rank takes part even with 0 keys, as in the reference (:843,911-917)
    parallel_for(m->p, [&](int r) {
      added[r] = kh_insert_dev(m->table[r], x.d_recv[r].p, nullptr, x.recv_n[r], default_val, 1);
      if (added[r] < 0) throw CudaError{(int)added[r]};
    });
    int64_t tot = 0; for (auto a : added) tot += a;
    return tot;
  });
}

// (key, value) pairs (batched_robinhood_map under DiscardReducer / ReplaceReducer, benchmark/BenchmarkDistHashTables.cpp:976-1020;
// insert_p with pair payloads, KH/distributed_batched_robinhood_map.hpp:910-1179).  The value a key keeps depends on the
// order in which its owner sees the pairs: sender by sender in rank order, input order inside a sender -- which is the order
// of the receive buffers here (exact offsets per sender, STABLE scatter), so one ordered insert per GPU reproduces it.
int64_t khd_insert_pairs_host(khd_map *m, const uint64_t *keys, const uint32_t *vals, uint64_t n) {
  return guarded([&]() -> int64_t {
    if (!m || (n && (!keys || !vals))) { set_error("null argument"); return KH_ERR_INVALID; }
    const int p = m->p;
    Exchange x;
    run_exchange(m, keys, n, true, x);
    std::vector<DevBuf<uint32_t>> d_rv(p);
    parallel_for(p, [&](int r) {
      DeviceGuard dg(m->dev[r]);
      d_rv[r].alloc(std::max<uint64_t>(x.recv_n[r], 1));
    });
    parallel_for(p, [&](int g) {
      DeviceGuard dg(m->dev[g]);
      cudaStream_t s = m->stream[g];
      const uint64_t ng = x.slice_beg[g + 1] - x.slice_beg[g];
      if (ng == 0) return;
      DevBuf<uint32_t> v(ng), pv(ng);
      KH_CUDA(cudaMemcpyAsync(v.p, vals + x.slice_beg[g], ng * sizeof(uint32_t), cudaMemcpyHostToDevice, s));
      if (p == 1) {
        KH_CUDA(cudaMemcpyAsync(d_rv[0].p, v.p, ng * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
      } else {
        KH_LAUNCH(gather_vals_kernel, (int)std::min<uint64_t>((ng + 255) / 256, 1184), 256, 0, s, v.p, x.d_perm[g].p, ng, pv.p);
        for (int r = 0; r < p; ++r) if (x.C[g][r])
          KH_CUDA(cudaMemcpyPeerAsync(d_rv[r].p + x.recv_off[r][g], m->dev[r], pv.p + x.plan[g].seg_h[r], m->dev[g],
                                      x.C[g][r] * sizeof(uint32_t), s));
      }
      KH_CUDA(cudaStreamSynchronize(s));
    });
    std::vector<int64_t> added(p, 0);
    parallel_for(p, [&](int r) {
      added[r] = kh_insert_dev(m->table[r], x.d_recv[r].p, d_rv[r].p, x.recv_n[r], 0, 1);
      if (added[r] < 0) throw CudaError{(int)added[r]};
    });
    int64_t tot = 0; for (auto a : added) tot += a;
    return tot;
  });
}

int khd_set_load_factors(khd_map *m, double min_load, double max_load) {
  if (!m) return KH_ERR_INVALID;
  for (auto *t : m->table) { const int rc = kh_set_load_factors(t, min_load, max_load); if (rc != KH_OK) return rc; }
  return KH_OK;
}

int64_t khd_count_host(khd_map *m, const uint64_t *keys, uint64_t n, uint8_t *out) {
  return guarded([&]() -> int64_t {
    if (!m || (n && (!keys || !out))) { set_error("null argument"); return KH_ERR_INVALID; }
    Exchange x;
    run_exchange(m, keys, n, true, x);
    std::vector<DevBuf<uint8_t>> res(m->p);
    std::vector<int64_t> hit(m->p, 0);
    parallel_for(m->p, [&](int r) {
      DeviceGuard dg(m->dev[r]);
      res[r].alloc(std::max<uint64_t>(x.recv_n[r], 1));
      hit[r] = kh_count_dev(m->table[r], x.d_recv[r].p, x.recv_n[r], res[r].p);
      if (hit[r] < 0) throw CudaError{(int)hit[r]};
    });
    return_results<uint8_t>(m, x, res, out);
    int64_t tot = 0; for (auto a : hit) tot += a;
    return tot;
  });
}

int64_t khd_find_host(khd_map *m, const uint64_t *keys, uint64_t n, uint32_t *out, uint32_t nonexistent) {
  return guarded([&]() -> int64_t {
    if (!m || (n && (!keys || !out))) { set_error("null argument"); return KH_ERR_INVALID; }
    Exchange x;
    run_exchange(m, keys, n, true, x);
    std::vector<DevBuf<uint32_t>> res(m->p);
    std::vector<int64_t> hit(m->p, 0);
    parallel_for(m->p, [&](int r) {
      DeviceGuard dg(m->dev[r]);
      res[r].alloc(std::max<uint64_t>(x.recv_n[r], 1));
      hit[r] = kh_find_dev(m->table[r], x.d_recv[r].p, x.recv_n[r], res[r].p, nonexistent);
      if (hit[r] < 0) throw CudaError{(int)hit[r]};
    });
    return_results<uint32_t>(m, x, res, out);
    int64_t tot = 0; for (auto a : hit) tot += a;
    return tot;
  });
}

int64_t khd_erase_host(khd_map *m, const uint64_t *keys, uint64_t n) {
  return guarded([&]() -> int64_t {
    if (!m || (n && !keys)) { set_error("null argument"); return KH_ERR_INVALID; }
    Exchange x;
    run_exchange(m, keys, n, false, x);
    std::vector<int64_t> gone(m->p, 0);
    parallel_for(m->p, [&](int r) {
      gone[r] = kh_erase_dev(m->table[r], x.d_recv[r].p, x.recv_n[r]);
      if (gone[r] < 0) throw CudaError{(int)gone[r]};
    });
    int64_t tot = 0; for (auto a : gone) tot += a;
    return tot;
  });
}

uint64_t khd_size(const khd_map *m) {
  uint64_t s = 0;
  if (m) for (auto *t : m->table) s += kh_size(t);
  return s;
}
uint64_t khd_local_size(const khd_map *m, int rank) {
  return (m && rank >= 0 && rank < m->p) ? kh_size(m->table[rank]) : 0;
}
int64_t khd_export_rank_host(khd_map *m, int rank, uint64_t *out_keys, uint32_t *out_vals, uint64_t cap) {
  if (!m || rank < 0 || rank >= m->p) { set_error("bad rank"); return KH_ERR_INVALID; }
  return kh_export_host(m->table[rank], out_keys, out_vals, cap);
}

}  // extern "C"
