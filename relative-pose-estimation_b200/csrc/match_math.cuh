// match_math.cuh — per-element arithmetic of the Hamming matcher (host/device, see rpe_hd.cuh).
//
// hamming256: popcount(q xor t) over 256 bits.  On B200 POPC runs on the XU pipe at 16 lanes/clk/SM
// (a quarter of the ALU rate) and the plain 8 x (XOR, POPC) form is bound by it (ncu:
// sm__inst_executed_pipe_xu 90 %).  Three carry-save adders (LOP3 0x96 = xor3, 0xE8 = majority) on
// the ALU pipe fold the eight xor words into two words of weight 1 and three of weight 2, so only
// five POPCs are issued per distance; with kCsa4 a fourth adder folds the weight-2 words into one
// of weight 2 and one of weight 4 (four POPCs).  Exact integer arithmetic either way.
#pragma once
#include "rpe_hd.cuh"

namespace rpe {

RPE_HD uint32_t popc32(uint32_t v) {
#if defined(__CUDA_ARCH__)
  return (uint32_t)__popc(v);
#else
  return (uint32_t)__builtin_popcount(v);
#endif
}
RPE_HD uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
#else
  return a ^ b ^ c;
#endif
}
RPE_HD uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) {
#if defined(__CUDA_ARCH__)
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
#else
  return (a & b) | (a & c) | (b & c);
#endif
}

template <bool kCsa4>
RPE_HD uint32_t hamming256(const uint32_t q[8], uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3, uint32_t t4,
                           uint32_t t5, uint32_t t6, uint32_t t7) {
  const uint32_t x0 = q[0] ^ t0, x1 = q[1] ^ t1, x2 = q[2] ^ t2, x3 = q[3] ^ t3;
  const uint32_t x4 = q[4] ^ t4, x5 = q[5] ^ t5, x6 = q[6] ^ t6, x7 = q[7] ^ t7;
  const uint32_t s0 = xor3(x0, x1, x2), c0 = maj3(x0, x1, x2);
  const uint32_t s1 = xor3(x3, x4, x5), c1 = maj3(x3, x4, x5);
  const uint32_t s2 = xor3(s0, s1, x6), c2 = maj3(s0, s1, x6);
  const uint32_t ones = popc32(s2) + popc32(x7);
  if (kCsa4) {
    const uint32_t s3 = xor3(c0, c1, c2), c3 = maj3(c0, c1, c2);
    return ones + 2u * popc32(s3) + 4u * popc32(c3);
  }
  return ones + 2u * (popc32(c0) + popc32(c1) + popc32(c2));
}

// Packed nearest-neighbour key (dist << 20 | idx) of one distance, built with multiply-adds so that the
// shifts and adds issue on the FMA pipe (IMAD) instead of the ALU pipe, which the XOR / carry-save
// network already fills (ncu: pipe_alu 73 %, pipe_xu 77 % with the plain form).  idx < 2^20.
template <bool kCsa4>
RPE_HD uint32_t hamming256_key(const uint32_t q[8], uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3, uint32_t t4,
                               uint32_t t5, uint32_t t6, uint32_t t7, uint32_t idx, uint32_t m20) {
  // m20 = 2^20 passed at run time: a literal would be strength-reduced to LEA / SHF on the ALU pipe
  const uint32_t x0 = q[0] ^ t0, x1 = q[1] ^ t1, x2 = q[2] ^ t2, x3 = q[3] ^ t3;
  const uint32_t x4 = q[4] ^ t4, x5 = q[5] ^ t5, x6 = q[6] ^ t6, x7 = q[7] ^ t7;
  const uint32_t s0 = xor3(x0, x1, x2), c0 = maj3(x0, x1, x2);
  const uint32_t s1 = xor3(x3, x4, x5), c1 = maj3(x3, x4, x5);
  const uint32_t s2 = xor3(s0, s1, x6), c2 = maj3(s0, s1, x6);
  const uint32_t m21 = 2u * m20, m22 = 4u * m20;
  uint32_t key = popc32(s2) * m20 + idx;
  key = popc32(x7) * m20 + key;
  if (kCsa4) {
    const uint32_t s3 = xor3(c0, c1, c2), c3 = maj3(c0, c1, c2);
    key = popc32(s3) * m21 + key;
    key = popc32(c3) * m22 + key;
  } else {
    key = popc32(c0) * m21 + key;
    key = popc32(c1) * m21 + key;
    key = popc32(c2) * m21 + key;
  }
  return key;
}

// ---- linearly pre-combined operands ("mixed" layout) -------------------------------------------------
// XOR is linear, so any XOR-combination of the words of x = q ^ t equals the same combination of q's words
// XOR the same combination of t's.  hamming_mix() rewrites a descriptor d[0..7] in place as
//   m = { d0, d1, d0^d1^d2, d3, d4, d3^d4^d5, d0^..^d6, d7 }
// (an invertible map, done once per descriptor when it is loaded), after which the SUM outputs of the three
// carry-save adders are single XORs of mixed words and the CARRY outputs follow from two addends and the sum
// (maj(a, b, a^b^s) = LOP3 0xD4 of (a, b, s)): 13 LOP3 + 4 POPC per 256-bit distance instead of 16 + 4, and the
// words x2, x5, x6 are never formed.  Exact integer arithmetic.
RPE_HD uint32_t carry_from_sum(uint32_t a, uint32_t b, uint32_t s) {   // maj(a, b, c) given s = a ^ b ^ c
#if defined(__CUDA_ARCH__)
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xD4;" : "=r"(d) : "r"(a), "r"(b), "r"(s));
  return d;
#else
  const uint32_t c = a ^ b ^ s;
  return (a & b) | (a & c) | (b & c);
#endif
}
RPE_HD void hamming_mix(uint32_t d[8]) {
  const uint32_t m2 = d[0] ^ d[1] ^ d[2], m5 = d[3] ^ d[4] ^ d[5];
  d[6] = m2 ^ m5 ^ d[6];
  d[2] = m2;
  d[5] = m5;
}
// q and t0..t7 in the mixed layout; kCsa4 as above (false: 11 LOP3 + 5 POPC)
template <bool kCsa4>
RPE_HD uint32_t hamming256_mixed_key(const uint32_t q[8], uint32_t t0, uint32_t t1, uint32_t t2, uint32_t t3,
                                     uint32_t t4, uint32_t t5, uint32_t t6, uint32_t t7, uint32_t idx, uint32_t m20) {
  const uint32_t x0 = q[0] ^ t0, x1 = q[1] ^ t1, s0 = q[2] ^ t2;
  const uint32_t x3 = q[3] ^ t3, x4 = q[4] ^ t4, s1 = q[5] ^ t5;
  const uint32_t s2 = q[6] ^ t6, x7 = q[7] ^ t7;
  const uint32_t c0 = carry_from_sum(x0, x1, s0), c1 = carry_from_sum(x3, x4, s1), c2 = carry_from_sum(s0, s1, s2);
  const uint32_t m21 = 2u * m20, m22 = 4u * m20;
  uint32_t key = popc32(s2) * m20 + idx;
  key = popc32(x7) * m20 + key;
  if (kCsa4) {
    const uint32_t s3 = xor3(c0, c1, c2), c3 = maj3(c0, c1, c2);
    key = popc32(s3) * m21 + key;
    key = popc32(c3) * m22 + key;
  } else {
    key = popc32(c0) * m21 + key;
    key = popc32(c1) * m21 + key;
    key = popc32(c2) * m21 + key;
  }
  return key;
}

// reference form (tests): eight independent popcounts
RPE_HD uint32_t hamming256_plain(const uint32_t q[8], const uint32_t t[8]) {
  uint32_t d = 0;
  for (int i = 0; i < 8; ++i) d += popc32(q[i] ^ t[i]);
  return d;
}

}  // namespace rpe
