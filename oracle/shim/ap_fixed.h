// TEST INFRASTRUCTURE ONLY (oracle build).  Stand-in for Vivado-HLS `ap_fixed.h`
// (absent from /root/reference).  Signed fixed point with
// W total bits, I integer bits (F = W-I fractional bits), raw two's-complement
// storage of the same size as ap_int<W> so the reference's bit-casts work
// (foldedmv-offload.h:139-140, interpret.hpp:174-180, cnvW1A1/hw/top.cpp:84).
#pragma once
#include "ap_int.h"

template <int W, int I, ap_q_mode Q = AP_TRN, ap_o_mode O = AP_WRAP, int N = 0> struct ap_fixed {
  typedef typename ap_shim::store<W>::type store_t;
  store_t v;
  static const int width = W;
  static const int iwidth = I;
  static const int F = W - I;

  long long raw() const {
    if (W < 64) { uint64_t m = 1ull << (W - 1); return (long long)(((uint64_t)v ^ m) - m); }
    return (long long)(uint64_t)v;
  }
  // store a (possibly out-of-range) raw value with this type's overflow mode
  void set_raw(long long r) {
    if (O == AP_SAT && W < 64) {
      long long hi = (long long)((1ull << (W - 1)) - 1), lo = -hi - 1;
      if (r > hi) r = hi;
      if (r < lo) r = lo;
    }
    v = (store_t)((uint64_t)r & ap_shim::mask_n(W));
  }
  // re-quantise a raw value that has F2 fractional bits
  void from_raw(long long r, int F2) {
    if (F >= F2) { set_raw((long long)((uint64_t)r << (F - F2))); return; }
    int sh = F2 - F;
    if (Q == AP_RND) r += 1ll << (sh - 1);  // round half toward +inf
    set_raw(r >> sh);                       // arithmetic shift = floor (AP_TRN)
  }

  ap_fixed() : v(0) {}
  ap_fixed(int x) { from_raw((long long)x, 0); }
  ap_fixed(long long x) { from_raw(x, 0); }
  ap_fixed(double x) {
    double s = std::ldexp(x, F);
    double r = (Q == AP_RND) ? std::floor(s + 0.5) : std::floor(s);
    if (O == AP_SAT && W < 64) {
      double hi = std::ldexp(1.0, W - 1) - 1, lo = -std::ldexp(1.0, W - 1);
      if (r > hi) r = hi;
      if (r < lo) r = lo;
    }
    set_raw((long long)r);
  }
  ap_fixed(float x) : ap_fixed((double)x) {}
  template <int W2, int I2, ap_q_mode Q2, ap_o_mode O2, int N2> ap_fixed(const ap_fixed<W2, I2, Q2, O2, N2> &o) { from_raw(o.raw(), W2 - I2); }
  template <int W2, bool S2> ap_fixed(const ap_int_base<W2, S2, false> &o) { from_raw(o.sget(), 0); }

  template <int W2, int I2, ap_q_mode Q2, ap_o_mode O2, int N2> ap_fixed &operator+=(const ap_fixed<W2, I2, Q2, O2, N2> &o) {
    const int F2 = W2 - I2;
    if (F >= F2) set_raw(raw() + (long long)((uint64_t)o.raw() << (F - F2)));
    else { ap_fixed t(o); set_raw(raw() + t.raw()); }
    return *this;
  }
  ap_fixed<W + 1, I + 1, Q, O, N> operator-() const { ap_fixed<W + 1, I + 1, Q, O, N> r; r.set_raw(-raw()); return r; }
  bool operator<(const ap_fixed &o) const { return raw() < o.raw(); }
  bool operator>(const ap_fixed &o) const { return raw() > o.raw(); }
  bool operator==(const ap_fixed &o) const { return raw() == o.raw(); }
  double to_double() const { return std::ldexp((double)raw(), -F); }
  explicit operator double() const { return to_double(); }
};

// ap_int * ap_fixed (interpret.hpp:87 decltype, and the W2A2 layer-0 product, cnvW2A2/hw/top.cpp:215)
template <int W1, bool S1, int W2, int I2, ap_q_mode Q2, ap_o_mode O2, int N2>
ap_fixed<W1 + W2, W1 + I2> operator*(const ap_int_base<W1, S1, false> &a, const ap_fixed<W2, I2, Q2, O2, N2> &b) {
  ap_fixed<W1 + W2, W1 + I2> r;
  r.set_raw(a.sget() * b.raw());
  return r;
}
template <int W1, bool S1, int W2, int I2, ap_q_mode Q2, ap_o_mode O2, int N2>
ap_fixed<W1 + W2, W1 + I2> operator*(const ap_fixed<W2, I2, Q2, O2, N2> &b, const ap_int_base<W1, S1, false> &a) { return a * b; }

template <int W, bool S>
template <int W2, int I2, ap_q_mode Q2, ap_o_mode O2, int N2>
ap_int_base<W, S, false>::ap_int_base(const ap_fixed<W2, I2, Q2, O2, N2> &f) {
  long long r = f.raw();
  const int F2 = W2 - I2;
  long long q = F2 > 0 ? (r >> F2) : (long long)((uint64_t)r << (-F2));
  if (F2 > 0 && r < 0 && (r & (long long)ap_shim::mask_n(F2))) ++q;  // toward zero (C-native)
  v = (store_t)((uint64_t)q & MASK());
}
template <int W, int I, ap_q_mode Q, ap_o_mode O, int N> std::ostream &operator<<(std::ostream &o, const ap_fixed<W, I, Q, O, N> &a) { return o << a.to_double(); }
