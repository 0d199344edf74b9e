// kh_hash.cuh -- device hash functions, bit-exact to the reference's hash functors.
//
// One key per thread; nothing here is lane-parallel like the reference's AVX/SSE batch
// kernels (KH/murmurhash3_64_avx.hpp, murmurhash3_32_avx.hpp, murmurhash3_32_sse.hpp) --
// on a GPU the 32 lanes of a warp already are the "SIMD lanes".  Keys are 8 or 16 bytes
// (bliss::common::Kmer<K,DNA,uint64_t>, KI/common/kmer.hpp:115-170), so every function is
// specialised for W = 1 or 2 words and fully unrolled; the generic byte-tail code of the
// published algorithms is unreachable and omitted.
//
// Functor -> algorithm (KH/hash_new.hpp):
//   identity  :136-160   word 0 of the key
//   murmur32  :175-197   MurmurHash3_x86_32 (also murmur3avx32 / murmur3sse32)
//   murmur    :207-235   MurmurHash3_x64_128, h[0]
//   murmur_x86:242-265   MurmurHash3_x86_128, h1 | h2 << 32 (also murmur3avx64,
//                        test/unit/test_kmer_hash.cpp:512-520)
//   clhash    :279-304   CLHASH (ext/clhash/src/clhash.c:272-346), key = glibc rand() words
//   farm      :310-328   farmhashna::Hash64WithSeed (ext/kmerind/ext/farmhash/src/farmhash.cc:523-529)
//   farm32    :333-351   low 32 bits of farm on an SSE4.2+AES build (farmhash.cc:882-893,1437-1447)
//   crc32c    KH/crc32c_sse.hpp:131-167  crc32q chain from `seed`, no inversions
#pragma once
#include <cstdint>
#include "../../include/kmerhash_b200.h"

#define KH_HD __host__ __device__ __forceinline__

namespace kh {

struct HashParams {
  uint32_t seed;
  uint64_t cl0, cl1;      // clhash key words 0 and 1
  uint64_t cllen_lo, cllen_hi;  // clmul(len_bytes, key[132]) -- constant per key width
};

KH_HD constexpr int hash_bits(int kind) {
  return (kind == KH_HASH_MURMUR32 || kind == KH_HASH_MURMUR32AVX || kind == KH_HASH_MURMUR32SSE ||
          kind == KH_HASH_FARM32 || kind == KH_HASH_CRC32C) ? 32 : 64;   // MURMUR32FIN is rejected by every entry point (keys <= 4 bytes only)
}

KH_HD uint32_t rotl32(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }
KH_HD uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
KH_HD uint64_t rotr64(uint64_t x, int r) { return (x >> r) | (x << (64 - r)); }

KH_HD uint32_t fmix32(uint32_t h) {
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16; return h;
}
KH_HD uint64_t fmix64(uint64_t k) {
  k ^= k >> 33; k *= 0xff51afd7ed558ccdULL; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ULL; k ^= k >> 33; return k;
}

// ---- MurmurHash3_x86_32 over 8 / 16 bytes ------------------------------------------------
template <int W>
KH_HD uint32_t murmur3_x86_32(const uint64_t (&key)[W], uint32_t seed) {
  uint32_t h1 = seed;
#pragma unroll
  for (int i = 0; i < 2 * W; ++i) {
    uint32_t k1 = (uint32_t)(key[i >> 1] >> (32 * (i & 1)));
    k1 *= 0xcc9e2d51u; k1 = rotl32(k1, 15); k1 *= 0x1b873593u;
    h1 ^= k1; h1 = rotl32(h1, 13); h1 = h1 * 5 + 0xe6546b64u;
  }
  h1 ^= (uint32_t)(8 * W);
  return fmix32(h1);
}

// ---- MurmurHash3_x86_128, low 64 bits ----------------------------------------------------
template <int W>
KH_HD uint64_t murmur3_x86_128_lo64(const uint64_t (&key)[W], uint32_t seed) {
  const uint32_t c1 = 0x239b961bu, c2 = 0xab0e9789u, c3 = 0x38b34ae5u, c4 = 0xa1e38b93u;
  uint32_t h1 = seed, h2 = seed, h3 = seed, h4 = seed;
  if (W == 2) {  // one full 16-byte block
    uint32_t k1 = (uint32_t)key[0], k2 = (uint32_t)(key[0] >> 32);
    uint32_t k3 = (uint32_t)key[W - 1], k4 = (uint32_t)(key[W - 1] >> 32);
    k1 *= c1; k1 = rotl32(k1, 15); k1 *= c2; h1 ^= k1;
    h1 = rotl32(h1, 19); h1 += h2; h1 = h1 * 5 + 0x561ccd1bu;
    k2 *= c2; k2 = rotl32(k2, 16); k2 *= c3; h2 ^= k2;
    h2 = rotl32(h2, 17); h2 += h3; h2 = h2 * 5 + 0x0bcaa747u;
    k3 *= c3; k3 = rotl32(k3, 17); k3 *= c4; h3 ^= k3;
    h3 = rotl32(h3, 15); h3 += h4; h3 = h3 * 5 + 0x96cd1c35u;
    k4 *= c4; k4 = rotl32(k4, 18); k4 *= c1; h4 ^= k4;
    h4 = rotl32(h4, 13); h4 += h1; h4 = h4 * 5 + 0x32ac3b17u;
  } else {       // 8-byte tail: k2 then k1, no block mixing
    uint32_t k1 = (uint32_t)key[0], k2 = (uint32_t)(key[0] >> 32);
    k2 *= c2; k2 = rotl32(k2, 16); k2 *= c3; h2 ^= k2;
    k1 *= c1; k1 = rotl32(k1, 15); k1 *= c2; h1 ^= k1;
  }
  const uint32_t len = 8 * W;
  h1 ^= len; h2 ^= len; h3 ^= len; h4 ^= len;
  h1 += h2; h1 += h3; h1 += h4; h2 += h1; h3 += h1; h4 += h1;
  h1 = fmix32(h1); h2 = fmix32(h2); h3 = fmix32(h3); h4 = fmix32(h4);
  h1 += h2; h1 += h3; h1 += h4; h2 += h1;
  return (uint64_t)h1 | ((uint64_t)h2 << 32);
}

// ---- MurmurHash3_x64_128, h[0] -----------------------------------------------------------
template <int W>
KH_HD uint64_t murmur3_x64_128_h0(const uint64_t (&key)[W], uint32_t seed) {
  const uint64_t c1 = 0x87c37b91114253d5ULL, c2 = 0x4cf5ad432745937fULL;
  uint64_t h1 = seed, h2 = seed;
  if (W == 2) {
    uint64_t k1 = key[0], k2 = key[W - 1];
    k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
    h1 = rotl64(h1, 27); h1 += h2; h1 = h1 * 5 + 0x52dce729;
    k2 *= c2; k2 = rotl64(k2, 33); k2 *= c1; h2 ^= k2;
    h2 = rotl64(h2, 31); h2 += h1; h2 = h2 * 5 + 0x38495ab5;
  } else {
    uint64_t k1 = key[0];
    k1 *= c1; k1 = rotl64(k1, 31); k1 *= c2; h1 ^= k1;
  }
  const uint64_t len = 8 * W;
  h1 ^= len; h2 ^= len;
  h1 += h2; h2 += h1;
  h1 = fmix64(h1); h2 = fmix64(h2);
  h1 += h2;
  return h1;
}

// ---- farmhashna::Hash64WithSeed for len 8 / 16 (HashLen0to16 branch len >= 8) ------------
KH_HD uint64_t farm_len16_mul(uint64_t u, uint64_t v, uint64_t mul) {
  uint64_t a = (u ^ v) * mul; a ^= (a >> 47);
  uint64_t b = (v ^ a) * mul; b ^= (b >> 47); b *= mul; return b;
}
template <int W>
KH_HD uint64_t farm64(const uint64_t (&key)[W], uint64_t seed) {
  const uint64_t k2 = 0x9ae16a3b2f90404fULL;
  const uint64_t len = 8 * W;
  uint64_t mul = k2 + len * 2;
  uint64_t a = key[0] + k2;
  uint64_t b = key[W - 1];                 // Fetch(s + len - 8)
  uint64_t c = rotr64(b, 37) * mul + a;
  uint64_t d = (rotr64(a, 25) + b) * mul;
  uint64_t h = farm_len16_mul(c, d, mul);
  // Hash64WithSeeds(s, len, k2, seed) = HashLen16(Hash64(s,len) - k2, seed), Hash128to64 multiplier
  return farm_len16_mul(h - k2, seed, 0x9ddfea08eb382d69ULL);
}

// ---- CLHASH, short path (one 128-bit carry-less product) ---------------------------------
// Carry-less 32 x 32 -> 64 bit product with integer multipliers: the operands are cut into four classes of bits four
// positions apart, so that in an ordinary product of two classes at most 8 ones meet in any column -- the column sums
// (< 16) never carry into the next kept bit, and bit 0 of every 4-bit group is the XOR the carry-less product wants.
// 16 widening 32-bit multiplies (IMAD.WIDE) instead of 32 shift-and-xor steps.
KH_HD uint64_t clmul32(uint32_t x, uint32_t y) {
  const uint32_t x0 = x & 0x11111111u, x1 = x & 0x22222222u, x2 = x & 0x44444444u, x3 = x & 0x88888888u;
  const uint32_t y0 = y & 0x11111111u, y1 = y & 0x22222222u, y2 = y & 0x44444444u, y3 = y & 0x88888888u;
#define KH_M(a, b) ((uint64_t)(a) * (uint64_t)(b))
  uint64_t z0 = KH_M(x0, y0) ^ KH_M(x1, y3) ^ KH_M(x2, y2) ^ KH_M(x3, y1);
  uint64_t z1 = KH_M(x0, y1) ^ KH_M(x1, y0) ^ KH_M(x2, y3) ^ KH_M(x3, y2);
  uint64_t z2 = KH_M(x0, y2) ^ KH_M(x1, y1) ^ KH_M(x2, y0) ^ KH_M(x3, y3);
  uint64_t z3 = KH_M(x0, y3) ^ KH_M(x1, y2) ^ KH_M(x2, y1) ^ KH_M(x3, y0);
#undef KH_M
  return (z0 & 0x1111111111111111ULL) | (z1 & 0x2222222222222222ULL) | (z2 & 0x4444444444444444ULL) | (z3 & 0x8888888888888888ULL);
}
// 64 x 64 -> 128 (PCLMULQDQ): Karatsuba over GF(2)[x], three 32-bit products
KH_HD void clmul64(uint64_t a, uint64_t b, uint64_t &lo, uint64_t &hi) {
  const uint32_t a0 = (uint32_t)a, a1 = (uint32_t)(a >> 32), b0 = (uint32_t)b, b1 = (uint32_t)(b >> 32);
  const uint64_t z0 = clmul32(a0, b0), z2 = clmul32(a1, b1);
  const uint64_t z1 = clmul32(a0 ^ a1, b0 ^ b1) ^ z0 ^ z2;
  lo = z0 ^ (z1 << 32);
  hi = z2 ^ (z1 >> 32);
}
template <int W>
KH_HD uint64_t clhash64(const uint64_t (&key)[W], const HashParams &p) {
  uint64_t lo, hi;
  if (W == 2) clmul64(p.cl0 ^ key[0], p.cl1 ^ key[W - 1], lo, hi);
  else        clmul64(p.cl0 ^ key[0], p.cl1, lo, hi);
  lo ^= p.cllen_lo; hi ^= p.cllen_hi;
  // reduce modulo x^64 + x^4 + x^3 + x + 1 (clhash.c:91-106): Q2 = hi clmul 0x1B (128 bit),
  // Q3 = (Q2 >> 64) clmul 0x1B folded once more (the pshufb table), result = Q3 ^ lo(Q2) ^ lo
  uint64_t q2lo = hi ^ (hi << 1) ^ (hi << 3) ^ (hi << 4);
  uint64_t q2hi = (hi >> 63) ^ (hi >> 61) ^ (hi >> 60);
  uint64_t q3 = q2hi ^ (q2hi << 1) ^ (q2hi << 3) ^ (q2hi << 4);
  return q3 ^ q2lo ^ lo;
}

// ---- CRC32-C chained over the key bytes --------------------------------------------------
KH_HD uint32_t crc32c_u64(uint32_t crc, uint64_t w) {
  uint64_t c = (uint64_t)crc ^ w;   // crc32q: the 8 bytes enter LSB first
#pragma unroll 16
  for (int i = 0; i < 64; ++i) c = (c >> 1) ^ (0x82F63B78ULL & (0 - (c & 1)));
  return (uint32_t)c;
}
template <int W>
KH_HD uint32_t crc32c(const uint64_t (&key)[W], uint32_t seed) {
  uint32_t crc = seed;
#pragma unroll
  for (int i = 0; i < W; ++i) crc = crc32c_u64(crc, key[i]);
  return crc;
}

// ---- dispatch ----------------------------------------------------------------------------
template <int KIND, int W>
KH_HD uint64_t hash_key(const uint64_t (&key)[W], const HashParams &p) {
  if (KIND == KH_HASH_IDEN) return key[0];
  if (KIND == KH_HASH_MURMUR64) return murmur3_x64_128_h0<W>(key, p.seed);
  if (KIND == KH_HASH_MURMUR64AVX || KIND == KH_HASH_MURMUR_X86) return murmur3_x86_128_lo64<W>(key, p.seed);
  if (KIND == KH_HASH_MURMUR32 || KIND == KH_HASH_MURMUR32AVX || KIND == KH_HASH_MURMUR32SSE)
    return murmur3_x86_32<W>(key, p.seed);
  if (KIND == KH_HASH_FARM) return farm64<W>(key, (uint64_t)p.seed);
  if (KIND == KH_HASH_FARM32) return (uint32_t)farm64<W>(key, (uint64_t)p.seed);
  if (KIND == KH_HASH_CLHASH) return clhash64<W>(key, p);
  if (KIND == KH_HASH_CRC32C) return crc32c<W>(key, p.seed);
  return 0;
}

// canonical kind ids actually instantiated (aliases folded)
KH_HD constexpr int fold_kind(int kind) {
  return (kind == KH_HASH_MURMUR_X86) ? KH_HASH_MURMUR64AVX
       : (kind == KH_HASH_MURMUR32AVX || kind == KH_HASH_MURMUR32SSE) ? KH_HASH_MURMUR32 : kind;
}

// host: fills HashParams (clhash key words from the glibc rand() sequence)
HashParams make_hash_params(int kind, uint32_t seed, int key_words);
bool hash_kind_supported(int kind);

}  // namespace kh

// Dispatch a runtime (kind, key_words) to a template instantiation.  FN is a macro/lambda body
// using KIND and W as compile-time constants.
#define KH_DISPATCH_HASH_W(kind, kw, ...)                                                     \
  do {                                                                                        \
    const int _k = kh::fold_kind(kind);                                                       \
    if ((kw) == 1) { constexpr int W = 1; KH_DISPATCH_KIND_(_k, __VA_ARGS__) }                 \
    else { constexpr int W = 2; KH_DISPATCH_KIND_(_k, __VA_ARGS__) }                           \
  } while (0)
#define KH_DISPATCH_KIND_(k, ...)                                                             \
  switch (k) {                                                                                \
    case KH_HASH_IDEN:        { constexpr int KIND = KH_HASH_IDEN; __VA_ARGS__; } break;      \
    case KH_HASH_MURMUR64:    { constexpr int KIND = KH_HASH_MURMUR64; __VA_ARGS__; } break;  \
    case KH_HASH_MURMUR64AVX: { constexpr int KIND = KH_HASH_MURMUR64AVX; __VA_ARGS__; } break; \
    case KH_HASH_MURMUR32:    { constexpr int KIND = KH_HASH_MURMUR32; __VA_ARGS__; } break;  \
    case KH_HASH_FARM:        { constexpr int KIND = KH_HASH_FARM; __VA_ARGS__; } break;      \
    case KH_HASH_FARM32:      { constexpr int KIND = KH_HASH_FARM32; __VA_ARGS__; } break;    \
    case KH_HASH_CLHASH:      { constexpr int KIND = KH_HASH_CLHASH; __VA_ARGS__; } break;    \
    case KH_HASH_CRC32C:      { constexpr int KIND = KH_HASH_CRC32C; __VA_ARGS__; } break;    \
    default: break;                                                                           \
  }
