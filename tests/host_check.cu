// tests/host_check.cu -- TEST INFRASTRUCTURE.  Compiles the product's __host__ __device__ arithmetic
// (kmerhash_b200/csrc/kh_hash.cuh, kh_kmer.cuh) for the HOST so that the CPU test tier can compare
// the very same source the kernels inline against the oracle, without a GPU.  Nothing in the
// product links or loads this file.
#include <cstdint>
#include <cstddef>
#include "../kmerhash_b200/csrc/kh_hash.cuh"
#include "../kmerhash_b200/csrc/kh_kmer.cuh"

using namespace kh;

template <int W>
static void hash_all(int kind, const HashParams &hp, const uint64_t *keys, size_t n, uint64_t *out) {
  for (size_t i = 0; i < n; ++i) {
    uint64_t key[W];
    for (int q = 0; q < W; ++q) key[q] = keys[i * W + q];
    uint64_t h = 0;
    switch (fold_kind(kind)) {
      case KH_HASH_IDEN:        h = hash_key<KH_HASH_IDEN, W>(key, hp); break;
      case KH_HASH_MURMUR64:    h = hash_key<KH_HASH_MURMUR64, W>(key, hp); break;
      case KH_HASH_MURMUR64AVX: h = hash_key<KH_HASH_MURMUR64AVX, W>(key, hp); break;
      case KH_HASH_MURMUR32:    h = hash_key<KH_HASH_MURMUR32, W>(key, hp); break;
      case KH_HASH_FARM:        h = hash_key<KH_HASH_FARM, W>(key, hp); break;
      case KH_HASH_FARM32:      h = hash_key<KH_HASH_FARM32, W>(key, hp); break;
      case KH_HASH_CLHASH:      h = hash_key<KH_HASH_CLHASH, W>(key, hp); break;
      case KH_HASH_CRC32C:      h = hash_key<KH_HASH_CRC32C, W>(key, hp); break;
    }
    out[i] = h;
  }
}

extern "C" {

int khtest_hash(int kind, uint32_t seed, int W, const uint64_t *keys, size_t n, uint64_t *out) {
  if (!hash_kind_supported(kind) || W < 1 || W > 2) return -1;
  const HashParams hp = make_hash_params(kind, seed, W);
  if (W == 1) hash_all<1>(kind, hp, keys, n, out); else hash_all<2>(kind, hp, keys, n, out);
  return 0;
}

int khtest_kmer(int k, int mode /*0 canonical, 1 revcomp*/, const uint64_t *in, size_t n, uint64_t *out) {
  const int W = kmer_words(k);
  for (size_t i = 0; i < n; ++i) {
    if (W == 1) {
      uint64_t a[1] = {in[i]}, r[1];
      if (mode == 0) kmer_canonical<1>(a, k, r); else kmer_revcomp<1>(a, k, r);
      out[i] = r[0];
    } else {
      uint64_t a[2] = {in[2 * i], in[2 * i + 1]}, r[2];
      if (mode == 0) kmer_canonical<2>(a, k, r); else kmer_revcomp<2>(a, k, r);
      out[2 * i] = r[0]; out[2 * i + 1] = r[1];
    }
  }
  return 0;
}

int khtest_dna_code(int c) { return (int)dna_code((uint8_t)c); }

// DNA5 / DNA16 k-mers (alphabet ids 5 / 16; 4 = DNA): reverse complement and canonical form as the kernels compute them
int khtest_kmer_alpha(int alpha, int k, int mode /*0 canonical, 1 revcomp*/, const uint64_t *in, size_t n, uint64_t *out) {
  const int W = kmer_words_alpha(alpha, k);
  if (W < 1 || W > 2) return -1;
  for (size_t i = 0; i < n; ++i) {
    if (W == 1) {
      uint64_t a[1] = {in[i]}, r[1];
      if (mode == 0) kmer_canonical_alpha<1>(a, alpha, k, r); else kmer_revcomp_alpha<1>(a, alpha, k, r);
      out[i] = r[0];
    } else {
      uint64_t a[2] = {in[2 * i], in[2 * i + 1]}, r[2];
      if (mode == 0) kmer_canonical_alpha<2>(a, alpha, k, r); else kmer_revcomp_alpha<2>(a, alpha, k, r);
      out[2 * i] = r[0]; out[2 * i + 1] = r[1];
    }
  }
  return 0;
}
int khtest_alpha_code(int alpha, int c) {
  return alpha == KH_ALPHA_DNA ? (int)alpha_code<2>((uint8_t)c) : alpha == KH_ALPHA_DNA5 ? (int)alpha_code<3>((uint8_t)c) : (int)alpha_code<4>((uint8_t)c);
}
int khtest_kmer_words_alpha(int alpha, int k) { return kmer_words_alpha(alpha, k); }

}
