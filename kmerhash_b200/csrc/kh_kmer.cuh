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

// ---- other alphabets: DNA5 = the reference's 3-bit DNA6 table, DNA16 = 4-bit IUPAC codes -------------------
// KI/common/alphabets.hpp:197-292 (gap 0, A 1, C 3, T 4, G 6, N/X 7, all else 2), :648-739 (. A C M G R S V T W Y H K D B N = 0..15,
// U = T, all else 15).  A k-mer holds bits * k bits, newest base lowest; up to two words here: k <= 42 (DNA5), k <= 32 (DNA16).
// Reverse complement of both = reversal of ALL bits of the value (KI/common/kmer.hpp:1476-1497): reversing a code's bits complements it.
constexpr int KH_ALPHA_DNA = 4, KH_ALPHA_DNA5 = 5, KH_ALPHA_DNA16 = 16;
KH_HD int alpha_bits(int alpha) { return alpha == KH_ALPHA_DNA ? 2 : alpha == KH_ALPHA_DNA5 ? 3 : alpha == KH_ALPHA_DNA16 ? 4 : 0; }
KH_HD int kmer_words_alpha(int alpha, int k) { const int b = alpha_bits(alpha); return b ? (b * k + 63) / 64 : 0; }
KH_HD uint32_t dna5_code(uint8_t c) {
  const uint8_t u = c & 0xDF;
  return (c == '-' || c == '.') ? 0u : (u == 'A') ? 1u : (u == 'C') ? 3u : (u == 'T') ? 4u : (u == 'G') ? 6u : (u == 'N' || u == 'X') ? 7u : 2u;
}
KH_HD uint32_t dna16_code(uint8_t c) {
  if (c == '-' || c == '.') return 0u;
  switch (c & 0xDF) {
    case 'A': return 1u;  case 'C': return 2u;  case 'M': return 3u;  case 'G': return 4u;  case 'R': return 5u;
    case 'S': return 6u;  case 'V': return 7u;  case 'T': case 'U': return 8u;  case 'W': return 9u;  case 'Y': return 10u;
    case 'H': return 11u; case 'K': return 12u; case 'D': return 13u; case 'B': return 14u;
    default: return 15u;
  }
}
template <int B> KH_HD uint32_t alpha_code(uint8_t c) { return B == 2 ? dna_code(c) : B == 3 ? dna5_code(c) : dna16_code(c); }
// complement of one code: DNA 3 - c; DNA5 / DNA16: the code's bits reversed
template <int B> KH_HD uint32_t alpha_comp(uint32_t c) {
  if (B == 2) return 3u - c;
  if (B == 3) return ((c & 1u) << 2) | (c & 2u) | (c >> 2);
  return ((c & 1u) << 3) | ((c & 2u) << 1) | ((c & 4u) >> 1) | (c >> 3);
}
KH_HD uint64_t brev64(uint64_t x) {
#ifdef __CUDA_ARCH__
  return ((uint64_t)__brev((uint32_t)x) << 32) | __brev((uint32_t)(x >> 32));
#else
  x = ((x >> 1) & 0x5555555555555555ULL) | ((x & 0x5555555555555555ULL) << 1);
  x = ((x >> 2) & 0x3333333333333333ULL) | ((x & 0x3333333333333333ULL) << 2);
  x = ((x >> 4) & 0x0F0F0F0F0F0F0F0FULL) | ((x & 0x0F0F0F0F0F0F0F0FULL) << 4);
  x = ((x >> 8) & 0x00FF00FF00FF00FFULL) | ((x & 0x00FF00FF00FF00FFULL) << 8);
  x = ((x >> 16) & 0x0000FFFF0000FFFFULL) | ((x & 0x0000FFFF0000FFFFULL) << 16);
  return (x >> 32) | (x << 32);
#endif
}
template <int W>
KH_HD void kmer_revcomp_alpha(const uint64_t (&in)[W], int alpha, int k, uint64_t (&out)[W]) {
  if (alpha == KH_ALPHA_DNA) { kmer_revcomp<W>(in, k, out); return; }
  const int sh = 64 * W - alpha_bits(alpha) * k;   // 0 <= sh < 64
  if (W == 1) { out[0] = brev64(in[0]) >> sh; return; }
  const uint64_t t0 = brev64(in[W - 1]), t1 = brev64(in[0]);
  out[0] = sh ? ((t0 >> sh) | (t1 << (64 - sh))) : t0;
  out[W - 1] = t1 >> sh;
}
template <int W>
KH_HD void kmer_canonical_alpha(const uint64_t (&in)[W], int alpha, int k, uint64_t (&out)[W]) {
  uint64_t rc[W];
  kmer_revcomp_alpha<W>(in, alpha, k, rc);
  const bool fwd = kmer_less<W>(in, rc);
#pragma unroll
  for (int i = 0; i < W; ++i) out[i] = fwd ? in[i] : rc[i];
}

}  // namespace kh
