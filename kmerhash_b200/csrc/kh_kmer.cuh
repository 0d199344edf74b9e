// kh_kmer.cuh -- 2-bit packed DNA k-mer arithmetic (k <= 64, 1 or 2 words).
//
// Layout = bliss::common::Kmer<K,DNA,uint64_t> (KI/common/kmer.hpp:115-170): word 0 least
// significant, newest base in bits [0,2) of word 0, unused top bits zero.
//   reverse complement (kmer.hpp:1078-1087,1453-1474): reverse the 2-bit groups of the whole
//     value, bitwise NOT, shift right by 64*W - 2k.
//   operator<  (kmer.hpp:775-845): whole-word compare, most significant word first.
//   lex_less   (KI/common/kmer_transform.hpp:109-125): min(kmer, revcomp).
//   DNA alphabet (KI/common/alphabets.hpp:139-161): A/a=0 C/c=1 G/g=2 T/t=3, everything else 0.
#pragma once
#include <cstdint>
#include "kh_hash.cuh"

namespace kh {

KH_HD int kmer_words(int k) { return (2 * k + 63) / 64; }

// reverse the order of the 32 2-bit groups of a word
KH_HD uint64_t rev2bit64(uint64_t x) {
#ifdef __CUDA_ARCH__
  // full bit reversal (BREV) then swap the two bits inside every group
  uint32_t lo = __brev((uint32_t)(x >> 32)), hi = __brev((uint32_t)x);
  uint64_t r = ((uint64_t)hi << 32) | lo;
  return ((r >> 1) & 0x5555555555555555ULL) | ((r & 0x5555555555555555ULL) << 1);
#else
  x = ((x >> 2) & 0x3333333333333333ULL) | ((x & 0x3333333333333333ULL) << 2);
  x = ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL) | ((x & 0x0F0F0F0F0F0F0F0FULL) << 4);
  x = ((x >> 8) & 0x00FF00FF00FF00FFULL) | ((x & 0x00FF00FF00FF00FFULL) << 8);
  x = ((x >> 16) & 0x0000FFFF0000FFFFULL) | ((x & 0x0000FFFF0000FFFFULL) << 16);
  return (x >> 32) | (x << 32);
#endif
}

template <int W>
KH_HD void kmer_revcomp(const uint64_t (&in)[W], int k, uint64_t (&out)[W]) {
  const int sh = 64 * W - 2 * k;   // 0 <= sh < 64
  if (W == 1) {
    uint64_t t = ~rev2bit64(in[0]);
    out[0] = t >> sh;
  } else {
    uint64_t t0 = ~rev2bit64(in[W - 1]);   // becomes the low word
    uint64_t t1 = ~rev2bit64(in[0]);       // becomes the high word
    out[0] = sh ? ((t0 >> sh) | (t1 << (64 - sh))) : t0;
    out[W - 1] = t1 >> sh;
  }
}

template <int W>
KH_HD bool kmer_less(const uint64_t (&a)[W], const uint64_t (&b)[W]) {
  if (W == 2) { if (a[W - 1] != b[W - 1]) return a[W - 1] < b[W - 1]; }
  return a[0] < b[0];
}

template <int W>
KH_HD void kmer_canonical(const uint64_t (&in)[W], int k, uint64_t (&out)[W]) {
  uint64_t rc[W];
  kmer_revcomp<W>(in, k, rc);
  const bool fwd = kmer_less<W>(in, rc);
#pragma unroll
  for (int i = 0; i < W; ++i) out[i] = fwd ? in[i] : rc[i];
}

// DNA::FROM_ASCII without a table: the low bits of the ASCII codes already separate ACGT.
//   'A'=0x41 'C'=0x43 'G'=0x47 'T'=0x54 (and lower case +0x20)
KH_HD uint32_t dna_code(uint8_t c) {
  uint8_t u = c & 0xDF;                    // fold case
  return (u == 'C') ? 1u : (u == 'G') ? 2u : (u == 'T') ? 3u : 0u;
}

}  // namespace kh
