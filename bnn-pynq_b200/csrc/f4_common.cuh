// e2m1 (fp4) helpers shared by the tcgen05 kind::mxf4 kernels.
// Codes used: 0 -> 0x0, +1 -> 0x2, -1 -> 0xA (and, for 2-bit weights, +-2 -> 0x4 / 0xC).  Even element in the low nibble.
//
// Inter-layer activation formats (fp4 core; ActFmt below; 2-bit activations: two bit planes nz, pos in ACT_BITS, codes {0x0, 0x2, 0xA} else):
//   ACT_BITS   uint32 [img][pixel][C/32]                 bit b of word j = channel 32j+b      (all cores; the tap / POPC format)
//   ACT_F4P    uint4  [img][plane C/32][pixel]           16 bytes = 32 channels as e2m1: exactly one row of a UMMA operand plane,
//                                                        so a consumer conv stages an image plane with ONE bulk copy, no expansion
//   ACT_F4T    uint4  [img/128][chunk K/32][img%128]     the same, blocked per 128-image tile in the operand layout of gemm_f4
//                                                        (chunk = pixel * C/32 + plane): a tile is one contiguous block
#pragma once
#include <climits>
#include <cstdint>

namespace bnnb {

enum ActFmt : int { ACT_BITS = 0, ACT_F4P = 1, ACT_F4T = 2, ACT_F4I = 3, ACT_F4PM = 4 };  // ACT_F4PM (conv_f4 input only): an ACT_F4P tensor
                                                                    // staged in interleaved row order by a 4-D tiled TMA load (tensor map)

// ACT_F4I: ACT_F4P with the images of a group interleaved pixel by pixel: uint4 [img/G][plane][pixel][img%G], G = the CONSUMER's
// images per group.  A consumer conv then stages a whole group's plane with one bulk copy AND its operand rows are ordered
// (pixel, image): the rows whose outputs can be valid - pixel index up to (OD-1)*IDIM + OD - form a prefix, so the tiles behind it
// (the bottom rows of every image, 14-45 % of the MMAs of these layers) are never computed.
__host__ __device__ inline uint32_t f4i_magic(int g) { return (uint32_t)((1ull << 32) / (uint32_t)g) + 1u; }  // img / g = umulhi(img, magic), img < 2^32 / g
__device__ __forceinline__ size_t f4i_index(uint32_t img, int plane, int px, int nplanes, int npx, int g, uint32_t magic) {
  const uint32_t grp = __umulhi(img, magic), il = img - grp * (uint32_t)g;
  return (((size_t)grp * nplanes + plane) * npx + px) * g + il;
}

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

// 32 two-bit channels (planes nz, pos) -> 16 bytes of e2m1 {-1, 0, +1}
__device__ __forceinline__ uint4 f4_from_tern(uint32_t nz, uint32_t pos) {
  const uint32_t ng = nz & ~pos;
  return make_uint4((f4_spread8(nz & 0xff) << 1) | (f4_spread8(ng & 0xff) << 3), (f4_spread8((nz >> 8) & 0xff) << 1) | (f4_spread8((ng >> 8) & 0xff) << 3),
                    (f4_spread8((nz >> 16) & 0xff) << 1) | (f4_spread8((ng >> 16) & 0xff) << 3), (f4_spread8(nz >> 24) << 1) | (f4_spread8(ng >> 24) << 3));
}

// ---- 2-bit outputs with the first threshold carried in the accumulator (conv_f4 / gemm_f4; cnv01_tc.cu has its own variant) -----
// Pre-load: 1.5 * 2^23 + 32767 - thr0, a float whose low mantissa bits are an integer; the MMAs add the exact sum S, so the low 16 bits
// of the accumulator are U = S - thr0 - 1 + 32768 and bit 15 of U is the decision thr0 < S.  The tensor core's fp32 accumulation is
// exact down to the last mantissa bit at this magnitude (measured on the fused kernel).  The epilogue reads two channels
// per register (tcgen05.ld .pack::16b: channel 2j in the low half of register j, 2j+1 in the high half), and one packed add of
// (thr0 - thr1) per register gives the same for the second threshold: thresholds are clamped to [-smax - 1, smax] with smax >= |S|
// (outside it they decide like the nearest bound), so no half ever leaves [0, 65536) and nothing carries between the halves.
constexpr float kTernBias = 12582912.0f + 32767.0f;
__device__ __forceinline__ int tern_clamp(int thr, int smax) { return thr == INT_MAX ? smax : max(min(thr, smax), -smax - 1); }
__device__ __forceinline__ float tern_preload(int thr0, int smax) { return kTernBias - (float)tern_clamp(thr0, smax); }
// (thr0 - thr1) of channels a (low half) and b (high half)
__device__ __forceinline__ uint32_t tern_delta(int thr0a, int thr1a, int thr0b, int thr1b, int smax) {
  return (uint32_t)(tern_clamp(thr0b, smax) - tern_clamp(thr1b, smax)) * 65536u + (uint32_t)(tern_clamp(thr0a, smax) - tern_clamp(thr1a, smax));
}
// bytes (a.1, b.1, a.3, b.3), each replaced by its sign bit replicated: four 16-bit decisions -> four byte masks, channel order
// (a.low, b.low, a.high, b.high)
__device__ __forceinline__ uint32_t tern_sign_bytes(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, 0xFBD9;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
// nibble masks of one decision for eight consecutive channels (registers 4q .. 4q+3), channel e in nibble e
__device__ __forceinline__ uint32_t tern_nibble_mask(uint32_t h0, uint32_t h1, uint32_t h2, uint32_t h3) {
  const uint32_t r1 = tern_sign_bytes(h0, h1), r2 = tern_sign_bytes(h2, h3);  // channels (0, 2, 1, 3), (4, 6, 5, 7)
  return (__byte_perm(r1, r2, 0x5410) & 0x0F0F0F0Fu) | (__byte_perm(r1, r2, 0x7632) & 0xF0F0F0F0u);
}
// 32 channels -> 16 bytes of e2m1: both thresholds exceeded -> 0x2 (+1), neither -> 0xA (-1), one -> 0x0
__device__ __forceinline__ uint4 tern_codes16(const uint32_t (&h)[16], const uint32_t (&d)[16]) {
  uint32_t v[4];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const uint32_t m0 = tern_nibble_mask(h[4 * q], h[4 * q + 1], h[4 * q + 2], h[4 * q + 3]);
    const uint32_t m1 = tern_nibble_mask(h[4 * q] + d[4 * q], h[4 * q + 1] + d[4 * q + 1], h[4 * q + 2] + d[4 * q + 2], h[4 * q + 3] + d[4 * q + 3]);
    v[q] = (~(m0 ^ m1) & 0xAAAAAAAAu) & (~m0 | 0x22222222u);
  }
  return make_uint4(v[0], v[1], v[2], v[3]);
}
// 1-bit outputs, 32 channels -> 16 bytes of e2m1 {0, +1}
__device__ __forceinline__ uint4 bin_codes16(const uint32_t (&h)[16]) {
  uint32_t v[4];
#pragma unroll
  for (int q = 0; q < 4; q++) v[q] = tern_nibble_mask(h[4 * q], h[4 * q + 1], h[4 * q + 2], h[4 * q + 3]) & 0x22222222u;
  return make_uint4(v[0], v[1], v[2], v[3]);
}
// one decision of 32 channels as a bit word (channel c at bit c)
__device__ __forceinline__ uint32_t tern_bit_word(const uint32_t (&y)[16]) {
  uint32_t lo = 0, hi = 0;
#pragma unroll
  for (int g = 0; g < 4; g++) {  // bytes: channels 2g, 8+2g, 2g+1, 8+2g+1 (+16 for hi); the odd channels land 16 bits too high
    const uint32_t sel = (0x0101u << (2 * g)) | (0x01010000u << (2 * g + 1));
    lo |= tern_sign_bytes(y[g], y[4 + g]) & sel;
    hi |= tern_sign_bytes(y[8 + g], y[12 + g]) & sel;
  }
  lo |= lo >> 16;
  hi |= hi >> 16;
  return __byte_perm(lo, hi, 0x5410);
}
// the two planes of ACT_BITS: nz = NOT(t0 XOR t1), pos = t0 AND t1
__device__ __forceinline__ void tern_planes16(const uint32_t (&h)[16], const uint32_t (&d)[16], uint32_t& nz, uint32_t& pos) {
  uint32_t y[16];
#pragma unroll
  for (int j = 0; j < 16; j++) y[j] = h[j] + d[j];
  const uint32_t t0 = tern_bit_word(h), t1 = tern_bit_word(y);
  nz = ~(t0 ^ t1);
  pos = t0 & t1;
}

}  // namespace bnnb
