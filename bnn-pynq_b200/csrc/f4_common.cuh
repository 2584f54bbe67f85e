// e2m1 (fp4) helpers shared by the tcgen05 kind::mxf4 kernels.
// Codes used: 0 -> 0x0, +1 -> 0x2, -1 -> 0xA (and, for 2-bit weights, +-2 -> 0x4 / 0xC).  Even element in the low nibble.
//
// Inter-layer activation formats (1-bit networks on the fp4 core; ActFmt below):
//   ACT_BITS   uint32 [img][pixel][C/32]                 bit b of word j = channel 32j+b      (all cores; the tap / POPC format)
//   ACT_F4P    uint4  [img][plane C/32][pixel]           16 bytes = 32 channels as e2m1: exactly one row of a UMMA operand plane,
//                                                        so a consumer conv stages an image plane with ONE bulk copy, no expansion
//   ACT_F4T    uint4  [img/128][chunk K/32][img%128]     the same, blocked per 128-image tile in the operand layout of gemm_f4
//                                                        (chunk = pixel * C/32 + plane): a tile is one contiguous block
#pragma once
#include <cstdint>

namespace bnnb {

enum ActFmt : int { ACT_BITS = 0, ACT_F4P = 1, ACT_F4T = 2 };

// 8 bits -> 8 nibbles (bit i at position 4i)
__device__ __forceinline__ uint32_t f4_spread8(uint32_t x) {
  x = (x | (x << 12)) & 0x000f000fu;
  x = (x | (x << 6)) & 0x03030303u;
  x = (x | (x << 3)) & 0x11111111u;
  return x;
}
// 32 one-bit channels -> 16 bytes of e2m1 {0, +1}
__device__ __forceinline__ uint4 f4_from_bits(uint32_t w) {
  return make_uint4(f4_spread8(w & 0xff) << 1, f4_spread8((w >> 8) & 0xff) << 1, f4_spread8((w >> 16) & 0xff) << 1, f4_spread8(w >> 24) << 1);
}
// 32 accumulators with folded thresholds (sign bit set <=> decision 0) -> 16 bytes of e2m1 {0, +1}: the top nibble of each fp32 /
// int32 word is shifted in (one funnel shift per channel), its sign bit is kept, inverted and moved to the code's bit 1
__device__ __forceinline__ uint4 f4_from_signs(const uint32_t (&r)[32]) {
  uint32_t v[4];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    uint32_t w = 0;
#pragma unroll
    for (int e = 7; e >= 0; e--) w = __funnelshift_l(r[q * 8 + e], w, 4);
    v[q] = (~w & 0x88888888u) >> 2;
  }
  return make_uint4(v[0], v[1], v[2], v[3]);
}

// Ternary decisions of 32 fp32 accumulators against two thresholds each: plane nz = NOT(t0 XOR t1), plane pos = t0 AND t1 with
// t = (thr < acc).  thr - acc is negative exactly when thr < acc (x - x = +0), so each decision is a sign bit: two FADDs (FMA pipe)
// and two funnel shifts (ALU pipe) per channel instead of two compares plus predicate-to-bit logic on the ALU pipe alone - the
// ternary epilogues are ALU-bound.  th0 / th1: 32 consecutive floats each (shared memory).
__device__ __forceinline__ void tern_planes(const uint32_t (&r)[32], const float* th0, const float* th1, uint32_t& nz, uint32_t& pos) {
  uint32_t u0[4] = {0, 0, 0, 0}, u1[4] = {0, 0, 0, 0};  // four independent chains each (a lone chain is pure latency)
#pragma unroll
  for (int q = 0; q < 4; q++) {
#pragma unroll
    for (int e = 7; e >= 0; e--) {
      const float s = __uint_as_float(r[8 * q + e]);
      u0[q] = __funnelshift_l(__float_as_uint(th0[8 * q + e] - s), u0[q], 1);
      u1[q] = __funnelshift_l(__float_as_uint(th1[8 * q + e] - s), u1[q], 1);
    }
  }
  const uint32_t t0 = __byte_perm(u0[0], u0[1], 0x1140) | __byte_perm(u0[2], u0[3], 0x4011);
  const uint32_t t1 = __byte_perm(u1[0], u1[1], 0x1140) | __byte_perm(u1[2], u1[3], 0x4011);
  nz = ~(t0 ^ t1);
  pos = t0 & t1;
}

}  // namespace bnnb
