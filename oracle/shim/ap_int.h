// TEST INFRASTRUCTURE ONLY (oracle build).  Stand-in for the Vivado-HLS header
// `ap_int.h`, which is not in /root/reference and not on this box
// (SURVEY.md §8c, Appendix D).  Written from the published arbitrary-precision
// integer semantics (two's-complement truncate / sign- or zero-extend); it lets
// the UNMODIFIED reference sources compile with g++ into oracle/_ref/.
// Nothing in the product path includes this file.
//
// Layout contract (hard; the reference host code reinterprets memory):
//   sizeof(ap_int<16>)==2  (foldedmv-offload.h:121-127 indexes ap_int<16>[])
//   sizeof(ap_uint<64>)==8 (foldedmv-offload.h:391 casts ExtMemWord*)
//   ap_uint<8> <-> ap_fixed<8,1> bit-cast (foldedmv-offload.h:140)
#pragma once
#include <cstdint>
#include <cstring>
#include <type_traits>
#include <iostream>
#include <cmath>

enum ap_q_mode { AP_RND, AP_RND_ZERO, AP_RND_MIN_INF, AP_RND_INF, AP_RND_CONV, AP_TRN, AP_TRN_ZERO };
enum ap_o_mode { AP_SAT, AP_SAT_ZERO, AP_SAT_SYM, AP_WRAP, AP_WRAP_SM };

namespace ap_shim {
template <int W, bool Small = (W <= 8), bool Half = (W <= 16), bool Word = (W <= 32)> struct store { typedef uint64_t type; };
template <int W> struct store<W, true, true, true> { typedef uint8_t type; };
template <int W> struct store<W, false, true, true> { typedef uint16_t type; };
template <int W> struct store<W, false, false, true> { typedef uint32_t type; };
inline uint64_t mask_n(int n) { return n >= 64 ? ~0ull : ((1ull << n) - 1); }
}  // namespace ap_shim

template <int W, bool S, bool Wide = (W > 64)> struct ap_int_base;
template <int W, int I, ap_q_mode Q, ap_o_mode O, int N> struct ap_fixed;

// ---------------------------------------------------------------- bit / range proxies
template <typename P> struct ap_bit_ref {
  P &p; int i;
  ap_bit_ref(P &p_, int i_) : p(p_), i(i_) {}
  operator bool() const { return p.get_bits(i, 1) != 0; }
  ap_bit_ref &operator=(unsigned long long x) { p.set_bits(i, 1, x & 1); return *this; }
  ap_bit_ref &operator=(const ap_bit_ref &o) { p.set_bits(i, 1, (bool)o); return *this; }
};

template <typename P> struct ap_range_ref {
  P &p; int hi, lo;
  ap_range_ref(P &p_, int hi_, int lo_) : p(p_), hi(hi_), lo(lo_) {}
  int len() const { return hi - lo + 1; }
  // chunk c (64 bits) of the range, zero-extended
  uint64_t chunk(int c) const {
    int off = c * 64, n = len() - off;
    if (n <= 0) return 0;
    return p.get_bits(lo + off, n > 64 ? 64 : n);
  }
  template <typename T> void assign_from(const T &src) {  // T has word(c) with sign/zero extension
    int n = len();
    for (int c = 0; c * 64 < n; ++c) {
      int m = n - c * 64; if (m > 64) m = 64;
      p.set_bits(lo + c * 64, m, src.word(c));
    }
  }
  template <int W2, bool S2, bool Wd2> ap_range_ref &operator=(const ap_int_base<W2, S2, Wd2> &o) { assign_from(o); return *this; }
  ap_range_ref &operator=(unsigned long long x) { ap_int_base<64, false, false> t(x); assign_from(t); return *this; }
  ap_range_ref &operator=(const ap_range_ref &o) {
    int n = len();
    for (int c = 0; c * 64 < n; ++c) { int m = n - c * 64; if (m > 64) m = 64; p.set_bits(lo + c * 64, m, o.chunk(c)); }
    return *this;
  }
  template <typename P2> ap_range_ref &operator=(const ap_range_ref<P2> &o) {
    int n = len();
    for (int c = 0; c * 64 < n; ++c) { int m = n - c * 64; if (m > 64) m = 64; p.set_bits(lo + c * 64, m, o.chunk(c)); }
    return *this;
  }
  operator unsigned long long() const { return chunk(0); }
};

// ---------------------------------------------------------------- W <= 64
template <int W, bool S> struct ap_int_base<W, S, false> {
  typedef typename ap_shim::store<W>::type store_t;
  typedef typename std::conditional<S, long long, unsigned long long>::type native_t;
  typedef ap_int_base self_t;
  store_t v;
  static const int width = W;
  static uint64_t MASK() { return ap_shim::mask_n(W); }

  ap_int_base() : v(0) {}
  template <typename T, typename std::enable_if<std::is_integral<T>::value, int>::type = 0>
  ap_int_base(T x) : v((store_t)((uint64_t)(long long)x & MASK())) {}
  ap_int_base(float x) : v((store_t)((uint64_t)(long long)x & MASK())) {}
  ap_int_base(double x) : v((store_t)((uint64_t)(long long)x & MASK())) {}
  template <int W2, bool S2, bool Wd2> ap_int_base(const ap_int_base<W2, S2, Wd2> &o) : v((store_t)(o.word(0) & MASK())) {}
  template <typename P> ap_int_base(const ap_range_ref<P> &r) : v((store_t)(r.chunk(0) & MASK())) {}
  template <typename P> ap_int_base(const ap_bit_ref<P> &b) : v((store_t)((bool)b ? 1 & MASK() : 0)) {}
  // value conversion from fixed point: integer part, C-native rounding (toward zero).
  // Real-header behaviour for negative non-integers is the SURVEY §8c residual (i).
  template <int W2, int I2, ap_q_mode Q2, ap_o_mode O2, int N2> ap_int_base(const ap_fixed<W2, I2, Q2, O2, N2> &f);

  long long sget() const {  // value, sign- or zero-extended
    if (S && W < 64) { uint64_t m = 1ull << (W - 1); return (long long)(((uint64_t)v ^ m) - m); }
    return (long long)(uint64_t)v;
  }
  uint64_t word(int c) const { return c == 0 ? (uint64_t)sget() : ((S && sget() < 0) ? ~0ull : 0ull); }
  operator native_t() const { return (native_t)sget(); }

  uint64_t get_bits(int lo, int n) const { return lo >= 64 ? 0 : (((uint64_t)v >> lo) & ap_shim::mask_n(n)); }
  void set_bits(int lo, int n, uint64_t x) {
    uint64_t m = ap_shim::mask_n(n) << lo;
    v = (store_t)((((uint64_t)v & ~m) | ((x << lo) & m)) & MASK());
  }
  bool operator[](int i) const { return ((uint64_t)v >> i) & 1; }
  ap_bit_ref<self_t> operator[](int i) { return ap_bit_ref<self_t>(*this, i); }
  ap_range_ref<self_t> operator()(int hi, int lo) { return ap_range_ref<self_t>(*this, hi, lo); }
  ap_range_ref<self_t> operator()(int hi, int lo) const { return ap_range_ref<self_t>(const_cast<self_t &>(*this), hi, lo); }
  ap_range_ref<self_t> range(int hi, int lo) { return (*this)(hi, lo); }

  template <typename T, typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0>
  self_t &operator+=(T x) { v = (store_t)((uint64_t)(sget() + (long long)x) & MASK()); return *this; }
  template <int W2, bool S2> self_t &operator+=(const ap_int_base<W2, S2, false> &o) { v = (store_t)((uint64_t)(sget() + o.sget()) & MASK()); return *this; }
  template <int W2, bool S2> self_t &operator|=(const ap_int_base<W2, S2, false> &o) { v = (store_t)(((uint64_t)v | o.word(0)) & MASK()); return *this; }
  self_t &operator++() { return *this += 1; }

  self_t operator>>(unsigned n) const { self_t r; r.v = (store_t)((uint64_t)(n >= 64 ? (sget() < 0 ? -1ll : 0ll) : (sget() >> n)) & MASK()); return r; }
  self_t operator<<(unsigned n) const { self_t r; r.v = (store_t)((n >= 64 ? 0ull : ((uint64_t)v << n)) & MASK()); return r; }
  self_t operator>>(int n) const { return *this >> (unsigned)n; }
  self_t operator<<(int n) const { return *this << (unsigned)n; }
  ap_int_base<W + 1, true, (W + 1 > 64)> operator-() const { return ap_int_base<W + 1, true, (W + 1 > 64)>(-sget()); }
};

#define AP_SHIM_BINOP(OP)                                                                                   \
  template <int W1, bool S1, int W2, bool S2>                                                               \
  ap_int_base<(W1 > W2 ? W1 : W2), (S1 && S2), false> operator OP(const ap_int_base<W1, S1, false> &a,       \
                                                                  const ap_int_base<W2, S2, false> &b) {    \
    return ap_int_base<(W1 > W2 ? W1 : W2), (S1 && S2), false>((unsigned long long)(a.word(0) OP b.word(0))); \
  }
AP_SHIM_BINOP(|)
AP_SHIM_BINOP(&)
AP_SHIM_BINOP(^)
#undef AP_SHIM_BINOP
#define AP_SHIM_CMP(OP)                                                                        \
  template <int W1, bool S1, int W2, bool S2>                                                  \
  bool operator OP(const ap_int_base<W1, S1, false> &a, const ap_int_base<W2, S2, false> &b) { \
    return a.sget() OP b.sget();                                                               \
  }
AP_SHIM_CMP(==)
AP_SHIM_CMP(!=)
AP_SHIM_CMP(<)
AP_SHIM_CMP(>)
AP_SHIM_CMP(<=)
AP_SHIM_CMP(>=)
#undef AP_SHIM_CMP
template <int W1, bool S1, int W2, bool S2>
ap_int_base<(W1 + W2 > 64 ? 64 : W1 + W2), true, false> operator*(const ap_int_base<W1, S1, false> &a, const ap_int_base<W2, S2, false> &b) {
  return ap_int_base<(W1 + W2 > 64 ? 64 : W1 + W2), true, false>(a.sget() * b.sget());
}
template <int W1, bool S1, int W2, bool S2>
ap_int_base<((W1 > W2 ? W1 : W2) + 1 > 64 ? 64 : (W1 > W2 ? W1 : W2) + 1), true, false> operator+(const ap_int_base<W1, S1, false> &a, const ap_int_base<W2, S2, false> &b) {
  return ap_int_base<((W1 > W2 ? W1 : W2) + 1 > 64 ? 64 : (W1 > W2 ? W1 : W2) + 1), true, false>(a.sget() + b.sget());
}
template <int W, bool S> std::ostream &operator<<(std::ostream &o, const ap_int_base<W, S, false> &a) { return o << a.sget(); }

// ---------------------------------------------------------------- W > 64 (stream words; unsigned use only)
template <int W, bool S> struct ap_int_base<W, S, true> {
  typedef ap_int_base self_t;
  static const int width = W;
  static const int NW = (W + 63) / 64;
  uint64_t v[NW];

  void norm() { if (W % 64) v[NW - 1] &= ap_shim::mask_n(W % 64); }
  ap_int_base() { for (int i = 0; i < NW; ++i) v[i] = 0; }
  template <typename T, typename std::enable_if<std::is_integral<T>::value, int>::type = 0>
  ap_int_base(T x) {
    v[0] = (uint64_t)(long long)x;
    uint64_t f = (std::is_signed<T>::value && (long long)x < 0) ? ~0ull : 0ull;
    for (int i = 1; i < NW; ++i) v[i] = f;
    norm();
  }
  template <int W2, bool S2, bool Wd2> ap_int_base(const ap_int_base<W2, S2, Wd2> &o) { for (int i = 0; i < NW; ++i) v[i] = o.word(i); norm(); }
  template <typename P> ap_int_base(const ap_range_ref<P> &r) { for (int i = 0; i < NW; ++i) v[i] = r.chunk(i); norm(); }

  uint64_t word(int c) const { return c < NW ? v[c] : 0ull; }
  uint64_t get_bits(int lo, int n) const {
    int w = lo >> 6, off = lo & 63;
    if (w >= NW) return 0;
    uint64_t r = v[w] >> off;
    if (off && w + 1 < NW) r |= v[w + 1] << (64 - off);
    return r & ap_shim::mask_n(n);
  }
  void set_bits(int lo, int n, uint64_t x) {
    int w = lo >> 6, off = lo & 63;
    uint64_t m = ap_shim::mask_n(n);
    x &= m;
    v[w] = (v[w] & ~(m << off)) | (x << off);
    if (off && off + n > 64 && w + 1 < NW) {
      int sh = 64 - off;
      v[w + 1] = (v[w + 1] & ~(m >> sh)) | (x >> sh);
    }
    norm();
  }
  bool operator[](int i) const { return (v[i >> 6] >> (i & 63)) & 1; }
  ap_bit_ref<self_t> operator[](int i) { return ap_bit_ref<self_t>(*this, i); }
  ap_range_ref<self_t> operator()(int hi, int lo) { return ap_range_ref<self_t>(*this, hi, lo); }
  ap_range_ref<self_t> operator()(int hi, int lo) const { return ap_range_ref<self_t>(const_cast<self_t &>(*this), hi, lo); }

  self_t operator>>(unsigned n) const {
    self_t r;
    int ws = n >> 6, bs = n & 63;
    for (int i = 0; i + ws < NW; ++i) {
      r.v[i] = v[i + ws] >> bs;
      if (bs && i + ws + 1 < NW) r.v[i] |= v[i + ws + 1] << (64 - bs);
    }
    return r;
  }
  self_t operator<<(unsigned n) const {
    self_t r;
    int ws = n >> 6, bs = n & 63;
    for (int i = NW - 1; i - ws >= 0; --i) {
      r.v[i] = v[i - ws] << bs;
      if (bs && i - ws - 1 >= 0) r.v[i] |= v[i - ws - 1] >> (64 - bs);
    }
    r.norm();
    return r;
  }
  self_t operator>>(int n) const { return *this >> (unsigned)n; }
  self_t operator<<(int n) const { return *this << (unsigned)n; }
  self_t operator|(const self_t &o) const { self_t r; for (int i = 0; i < NW; ++i) r.v[i] = v[i] | o.v[i]; return r; }
  self_t operator&(const self_t &o) const { self_t r; for (int i = 0; i < NW; ++i) r.v[i] = v[i] & o.v[i]; return r; }
  self_t &operator|=(const self_t &o) { for (int i = 0; i < NW; ++i) v[i] |= o.v[i]; return *this; }
  bool operator==(const self_t &o) const { for (int i = 0; i < NW; ++i) if (v[i] != o.v[i]) return false; return true; }
  bool operator!=(const self_t &o) const { return !(*this == o); }
  operator unsigned long long() const { return v[0]; }
};

// ---------------------------------------------------------------- user-facing names
template <int W> struct ap_int : ap_int_base<W, true> {
  typedef ap_int_base<W, true> B;
  ap_int() {}
  template <typename T> ap_int(const T &x) : B(x) {}
};
template <int W> struct ap_uint : ap_int_base<W, false> {
  typedef ap_int_base<W, false> B;
  ap_uint() {}
  template <typename T> ap_uint(const T &x) : B(x) {}
};
#include "ap_fixed.h"
