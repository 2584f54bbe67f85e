// TEST INFRASTRUCTURE ONLY (oracle build) -- never linked, imported or executed by the product path.
//
// Per-layer tap harness over the UNMODIFIED reference sources (oracle plan item 2).  This file is compiled
// once per network IN PLACE OF <net>/hw/top.cpp: it #includes that top.cpp, so the reference's own static parameter memories
// (weightsN / threshsN, cnvW1A1/hw/top.cpp:51-68) are loaded by the reference's own loader (load_parameters ->
// FoldedMVLoadLayerMem -> BlackBoxJam(doInit=1) -> DoMemInit), and then replays the DoCompute call list
// (cnvW1A1/hw/top.cpp:210-235, cnvW1A2/hw/top.cpp:209-235, cnvW2A2/hw/top.cpp:211-237, lfcW1A1/hw/top.cpp:155-164,
// lfcW1A2/hw/top.cpp:155-164) ONE LAYER AT A TIME with the reference's own templates and template arguments:
//   * every layer's input stream is duplicated; one copy feeds the layer exactly as DoCompute calls it (threshold activation),
//     the other feeds the SAME ConvLayer_Batch / StreamingFCLayer_Batch instantiation with PassThroughActivation
//     (activations.hpp:85-92) and a 16-bit output slice, which makes Matrix_Vector_Activate_Batch (mvau.hpp:86-172) emit its
//     accumulators (same accumulator type: decltype(activation.init()) is ap_int<16> in both, mvau.hpp:110);
//   * CNV layer 0 accumulates in ap_fixed<24,16>; its raw bits leave through a 4-line activation class that bit-casts them
//     (the inverse of interpret.hpp:174-180), since a range assignment from ap_fixed would convert by value;
//   * the stream after each layer (after the max-pool where one follows, maxpool.h:62-208) is dumped as the layer's output.
// Output format = oracle/bnn_oracle.c's taps: per image, concatenated over layers, int32 accumulators [ofm_dim^2][MatrixH] and
// int8 outputs [out_dim^2][MatrixH] (1-bit: 0/1; 2-bit: -1/0/+1; CNV last layer: the 16-bit score in acc, its low byte in act).
// tests/golden/make_golden_taps.py runs this in the build container and commits tests/golden/taps_<net>.npz.
#include "top.cpp"

#include <cstdint>
#include <vector>

namespace {

template <int W> std::vector<ap_uint<W>> drain(hls::stream<ap_uint<W>> &s) {
  std::vector<ap_uint<W>> v;
  while (!s.empty()) v.push_back(s.read());
  return v;
}
template <int W> void refill(hls::stream<ap_uint<W>> &s, const std::vector<ap_uint<W>> &v) {
  for (const auto &x : v) s.write(x);
}
template <int W> void tee(hls::stream<ap_uint<W>> &in, hls::stream<ap_uint<W>> &a, hls::stream<ap_uint<W>> &b) {
  auto v = drain(in);
  refill(a, v);
  refill(b, v);
}
// stream words -> fields of `bits` bits, LSB first (streamtools.h:255-315), sign- or zero-extended
template <int W> void fields(const std::vector<ap_uint<W>> &v, int bits, bool sgn, std::vector<int32_t> &out) {
  for (const auto &w : v)
    for (int lo = 0; lo + bits <= W; lo += bits) {
      uint32_t f = (uint32_t)(unsigned long long)ap_uint<32>(w(lo + bits - 1, lo));
      int32_t x = (int32_t)f;
      if (sgn && bits < 32 && (f >> (bits - 1))) x = (int32_t)(f | (~0u << bits));
      out.push_back(x);
    }
}

// raw bits of the ap_fixed<24,16> accumulator of CNV layer 0
struct RawFixedAccu {
  ap_fixed<24, 16> init(unsigned const, unsigned const) const { return ap_fixed<24, 16>(0); }
  ap_int<24> activate(unsigned const, unsigned const, ap_fixed<24, 16> const &accu) const {
    return *reinterpret_cast<ap_int<24> const *>(&accu);
  }
};

struct Sink {
  unsigned reps;
  int32_t *acc;
  int8_t *act;
  size_t acc_per_img, act_per_img, acc_off = 0, act_off = 0;
  // accumulator stream of one layer: [img][rows][mh] fields
  template <int W> void put_acc(hls::stream<ap_uint<W>> &s, int bits, size_t per_img) {
    std::vector<int32_t> f;
    fields(drain(s), bits, true, f);
    for (unsigned i = 0; i < reps; i++)
      for (size_t k = 0; k < per_img; k++) acc[i * acc_per_img + acc_off + k] = f[i * per_img + k];
    acc_off += per_img;
  }
  // output stream of one layer (kept intact for the next layer): [img][pixels][ch] fields
  template <int W> void put_act(hls::stream<ap_uint<W>> &s, int bits, bool sgn, size_t per_img) {
    auto v = drain(s);
    refill(s, v);
    std::vector<int32_t> f;
    fields(v, bits, sgn, f);
    for (unsigned i = 0; i < reps; i++)
      for (size_t k = 0; k < per_img; k++) act[i * act_per_img + act_off + k] = (int8_t)f[i * per_img + k];
    act_off += per_img;
  }
};

#define LUT ap_resource_lut()
#define CONV_GEOM(i) L##i##_K, L##i##_IFM_CH, L##i##_IFM_DIM, L##i##_OFM_CH, L##i##_OFM_DIM, L##i##_SIMD, L##i##_PE
#define FC_GEOM(i) L##i##_MW, L##i##_MH, L##i##_SIMD, L##i##_PE
}  // namespace

// the shim's stream has no value_type: name it through a trait
namespace {
template <typename S> struct stream_elem;
template <typename T> struct stream_elem<hls::stream<T>> { typedef T type; };
}  // namespace
// layer i exactly as DoCompute calls it (template arguments SRC, DST, WGT) + the accumulator replay
#define STREAM_LIKE(IN) hls::stream<typename stream_elem<typename std::remove_reference<decltype(IN)>::type>::type>
#define TAP_CONV(i, IN, OUT, SRC, DST, WGT)                                                                                     \
  {                                                                                                                             \
    STREAM_LIKE(IN) a_, b_;                                                                                                     \
    tee(IN, a_, b_);                                                                                                            \
    ConvLayer_Batch<CONV_GEOM(i), SRC, DST, WGT>(a_, OUT, weights##i, threshs##i, reps, LUT);                                   \
    hls::stream<ap_uint<L##i##_PE * 16>> q_;                                                                                    \
    ConvLayer_Batch<CONV_GEOM(i), SRC, Slice<ap_int<16>>, WGT>(b_, q_, weights##i, PassThroughActivation<ap_int<16>>(), reps, LUT); \
    sink.put_acc(q_, 16, (size_t)L##i##_OFM_DIM * L##i##_OFM_DIM * L##i##_OFM_CH);                                              \
  }
#define TAP_FC(i, IN, OUT, SRC, DST, WGT)                                                                                       \
  {                                                                                                                             \
    STREAM_LIKE(IN) a_, b_;                                                                                                     \
    tee(IN, a_, b_);                                                                                                            \
    StreamingFCLayer_Batch<FC_GEOM(i), SRC, DST, WGT>(a_, OUT, weights##i, threshs##i, reps, LUT);                              \
    hls::stream<ap_uint<L##i##_PE * 16>> q_;                                                                                    \
    StreamingFCLayer_Batch<FC_GEOM(i), SRC, Slice<ap_int<16>>, WGT>(b_, q_, weights##i, PassThroughActivation<ap_int<16>>(), reps, LUT); \
    sink.put_acc(q_, 16, (size_t)L##i##_MH);                                                                                    \
  }

#if defined(TAP_NET_cnvW1A1) || defined(TAP_NET_cnvW1A2) || defined(TAP_NET_cnvW2A2)
// ---- CNV: per-network template arguments, transcribed from the DoCompute call list of the network's own top.cpp ----------------
#if defined(TAP_NET_cnvW1A1)  // cnvW1A1/hw/top.cpp:214-233
#define A_BITS 1
#define SRC0 Slice<ap_fixed<8, 1, AP_TRN, AP_SAT>>
#define DSTX Identity
#define WGT0 Recast<Binary>
#define SRCX Recast<XnorMul>
#define WGTX Identity
#define SRC8 Recast<XnorMul>
#define DST8 Slice<ap_uint<16>>
#define POOL(DIM, CH, IN, OUT) StreamingMaxPool_Batch<DIM, 2, CH>(IN, OUT, reps)
#elif defined(TAP_NET_cnvW1A2)  // cnvW1A2/hw/top.cpp:213-232
#define A_BITS 2
#define SRC0 Slice<ap_fixed<8, 1, AP_TRN, AP_SAT>>
#define DSTX Slice<ap_int<2>>
#define WGT0 Recast<Binary>
#define SRCX Slice<ap_int<2>>
#define WGTX Recast<Binary>
#define SRC8 Slice<ap_int<2>>
#define DST8 Slice<ap_uint<16>>
#define POOL(DIM, CH, IN, OUT) StreamingMaxPool_Precision_Batch<DIM, 2, CH, ap_int<2>, -1>(IN, OUT, reps)
#else  // cnvW2A2/hw/top.cpp:215-234
#define A_BITS 2
#define SRC0 Slice<ap_fixed<8, 1, AP_TRN, AP_SAT>>
#define DSTX Slice<ap_int<2>>
#define WGT0 Identity
#define SRCX Slice<ap_int<2>>
#define WGTX Identity
#define SRC8 Slice<ap_int<2>>
#define DST8 Slice<ap_uint<16>>
#define POOL(DIM, CH, IN, OUT) StreamingMaxPool_Precision_Batch<DIM, 2, CH, ap_int<2>, -1>(IN, OUT, reps)
#endif

extern "C" size_t ref_tap_acc_size(void) {
  return (size_t)L0_OFM_DIM * L0_OFM_DIM * L0_OFM_CH + (size_t)L1_OFM_DIM * L1_OFM_DIM * L1_OFM_CH + (size_t)L2_OFM_DIM * L2_OFM_DIM * L2_OFM_CH +
         (size_t)L3_OFM_DIM * L3_OFM_DIM * L3_OFM_CH + (size_t)L4_OFM_DIM * L4_OFM_DIM * L4_OFM_CH + (size_t)L5_OFM_DIM * L5_OFM_DIM * L5_OFM_CH + L6_MH + L7_MH + L8_MH;
}
extern "C" size_t ref_tap_act_size(void) {
  return (size_t)L0_OFM_DIM * L0_OFM_DIM * L0_OFM_CH + (size_t)(L1_OFM_DIM / 2) * (L1_OFM_DIM / 2) * L1_OFM_CH + (size_t)L2_OFM_DIM * L2_OFM_DIM * L2_OFM_CH +
         (size_t)(L3_OFM_DIM / 2) * (L3_OFM_DIM / 2) * L3_OFM_CH + (size_t)L4_OFM_DIM * L4_OFM_DIM * L4_OFM_CH + (size_t)L5_OFM_DIM * L5_OFM_DIM * L5_OFM_CH + L6_MH + L7_MH + L8_MH;
}

// in: reps x 384 words (int8 HWC images as quantiseAndPack leaves them, foldedmv-offload.h:129-144); acc / act: reps x tap sizes
extern "C" int ref_tap(const unsigned long long *in, unsigned reps, int32_t *acc, int8_t *act) {
  Sink sink{reps, acc, act, ref_tap_acc_size(), ref_tap_act_size()};
  const bool sg = A_BITS == 2;
  hls::stream<ap_uint<64>> inter0, inter1, inter2, inter3, inter9, inter10, memOutStrm;
  hls::stream<ap_uint<192>> inter0_1;
  hls::stream<ap_uint<24>> inter0_2;
  hls::stream<ap_uint<128>> inter4, inter5, inter6;
  hls::stream<ap_uint<256>> inter7, inter8;
  Mem2Stream_Batch<64, 32 * 32 * 3>((ap_uint<64> *)in, inter0, reps);
  StreamingDataWidthConverter_Batch<64, 192, (32 * 32 * 3 * 8) / 64>(inter0, inter0_1, reps);
  StreamingDataWidthConverter_Batch<192, 24, (32 * 32 * 3 * 8) / 192>(inter0_1, inter0_2, reps);
  {  // layer 0: ap_fixed<24,16> accumulators
    hls::stream<ap_uint<24>> a_, b_;
    tee(inter0_2, a_, b_);
    ConvLayer_Batch<CONV_GEOM(0), SRC0, DSTX, WGT0>(a_, inter1, weights0, threshs0, reps, LUT);
    hls::stream<ap_uint<L0_PE * 24>> q_;
    ConvLayer_Batch<CONV_GEOM(0), SRC0, Slice<ap_int<24>>, WGT0>(b_, q_, weights0, RawFixedAccu(), reps, LUT);
    sink.put_acc(q_, 24, (size_t)L0_OFM_DIM * L0_OFM_DIM * L0_OFM_CH);
  }
  sink.put_act(inter1, A_BITS, sg, (size_t)L0_OFM_DIM * L0_OFM_DIM * L0_OFM_CH);
  TAP_CONV(1, inter1, inter2, SRCX, DSTX, WGTX)
  POOL(L1_OFM_DIM, L1_OFM_CH, inter2, inter3);
  sink.put_act(inter3, A_BITS, sg, (size_t)(L1_OFM_DIM / 2) * (L1_OFM_DIM / 2) * L1_OFM_CH);
  TAP_CONV(2, inter3, inter4, SRCX, DSTX, WGTX)
  sink.put_act(inter4, A_BITS, sg, (size_t)L2_OFM_DIM * L2_OFM_DIM * L2_OFM_CH);
  TAP_CONV(3, inter4, inter5, SRCX, DSTX, WGTX)
  POOL(L3_OFM_DIM, L3_OFM_CH, inter5, inter6);
  sink.put_act(inter6, A_BITS, sg, (size_t)(L3_OFM_DIM / 2) * (L3_OFM_DIM / 2) * L3_OFM_CH);
  TAP_CONV(4, inter6, inter7, SRCX, DSTX, WGTX)
  sink.put_act(inter7, A_BITS, sg, (size_t)L4_OFM_DIM * L4_OFM_DIM * L4_OFM_CH);
  TAP_CONV(5, inter7, inter8, SRCX, DSTX, WGTX)
  sink.put_act(inter8, A_BITS, sg, (size_t)L5_OFM_DIM * L5_OFM_DIM * L5_OFM_CH);
  TAP_FC(6, inter8, inter9, SRCX, DSTX, WGTX)
  sink.put_act(inter9, A_BITS, sg, (size_t)L6_MH);
  TAP_FC(7, inter9, inter10, SRCX, DSTX, WGTX)
  sink.put_act(inter10, A_BITS, sg, (size_t)L7_MH);
  {  // layer 8 has no thresholds: DoCompute itself passes the accumulators through (PassThroughActivation<ap_uint<16>>)
    WidthAdjustedOutputStream<16 * L8_PE, 64, L8_MH / L8_PE> wa_out(memOutStrm, reps);
    StreamingFCLayer_Batch<FC_GEOM(8), SRC8, DST8, WGTX>(inter10, static_cast<hls::stream<ap_uint<16 * L8_PE>> &>(wa_out), weights8,
                                                         PassThroughActivation<ap_uint<16>>(), reps, LUT);
  }
  {
    auto v = drain(memOutStrm);
    std::vector<int32_t> f;
    fields(v, 16, true, f);
    for (unsigned i = 0; i < reps; i++)
      for (int k = 0; k < L8_MH; k++) {
        acc[i * sink.acc_per_img + sink.acc_off + k] = f[(size_t)i * L8_MH + k];
        act[i * sink.act_per_img + sink.act_off + k] = (int8_t)f[(size_t)i * L8_MH + k];
      }
  }
  return 0;
}

#elif defined(TAP_NET_lfcW1A1) || defined(TAP_NET_lfcW1A2)
#if defined(TAP_NET_lfcW1A1)  // lfcW1A1/hw/top.cpp:156-163
#define A_BITS 1
#define SRC0 Recast<XnorMul>
#define DSTX Identity
#define WGTX Identity
#define SRCX Recast<XnorMul>
#define DST3 Identity
#else  // lfcW1A2/hw/top.cpp:156-163
#define A_BITS 2
#define SRC0 Recast<Binary>
#define DSTX Slice<ap_int<2>>
#define WGTX Recast<Binary>
#define SRCX Slice<ap_int<2>>
#define DST3 Slice<ap_int<1>>
#endif

extern "C" size_t ref_tap_acc_size(void) { return (size_t)L0_MH + L1_MH + L2_MH + L3_MH; }
extern "C" size_t ref_tap_act_size(void) { return (size_t)L0_MH + L1_MH + L2_MH + L3_MH; }

// in: reps x 13 words (832 bits per image as binarizeAndPack leaves them, foldedmv-offload.cpp:82-98)
extern "C" int ref_tap(const unsigned long long *in, unsigned reps, int32_t *acc, int8_t *act) {
  Sink sink{reps, acc, act, ref_tap_acc_size(), ref_tap_act_size()};
  const bool sg = A_BITS == 2;
  hls::stream<ap_uint<64>> memInStrm, memOutStrm;
  hls::stream<ap_uint<L0_PE * (L0_API + L0_APF)>> inter0;
  hls::stream<ap_uint<L1_PE * (L1_API + L1_APF)>> inter1;
  hls::stream<ap_uint<L2_PE * (L2_API + L2_APF)>> inter2;
  Mem2Stream_Batch<64, 832 / 8>((ap_uint<64> *)in, memInStrm, reps);
  TAP_FC(0, memInStrm, inter0, SRC0, DSTX, WGTX)
  sink.put_act(inter0, A_BITS, sg, (size_t)L0_MH);
  TAP_FC(1, inter0, inter1, SRCX, DSTX, WGTX)
  sink.put_act(inter1, A_BITS, sg, (size_t)L1_MH);
  TAP_FC(2, inter1, inter2, SRCX, DSTX, WGTX)
  sink.put_act(inter2, A_BITS, sg, (size_t)L2_MH);
  TAP_FC(3, inter2, memOutStrm, SRCX, DST3, WGTX)
  sink.put_act(memOutStrm, 1, false, (size_t)L3_MH);
  return 0;
}
#else
#error "compile with -DTAP_NET_<network>"
#endif
