// Network descriptors for the five FINN-style networks of BNN-PYNQ, transcribed from
// bnn/src/network/<net>/hw/config.h (folding: PE, SIMD, WMEM, TMEM, API, WPI) and
// bnn/src/network/<net>/hw/top.cpp (datatypes of every layer's MVAU).  Shared by host and device code.
#pragma once
#include <cstdint>

namespace bnnb {

enum LayerKind : int { KIND_CONV = 0, KIND_FC = 1 };
// how a layer reads its input vector (top.cpp template arguments TSrcI / TWeightI)
enum InMode : int {
  IN_FIX8 = 0,  // Slice<ap_fixed<8,1>>: CNV layer 0, int8 pixels          (cnvW1A1/hw/top.cpp:214)
  IN_XNOR = 1,  // Recast<XnorMul>: acc = #(w_bit == x_bit)                 (cnvW1A1/hw/top.cpp:215-233)
  IN_BIN = 2,   // Recast<Binary> x Recast<Binary>: signed +-1 sum          (lfcW1A2/hw/top.cpp:156)
  IN_TERN = 3,  // Slice<ap_int<2>>: x in {-1,0,+1}                         (cnvW1A2/hw/top.cpp:213-232)
};
enum OutMode : int {
  OUT_BIT = 0,    // ThresholdsActivation<..,1,..,ap_uint<1>>   -> 1 bit
  OUT_TERN = 1,   // ThresholdsActivation<..,2,..,ap_int<2>,-1> -> {-1,0,+1} as (nz,pos) planes
  OUT_SCORE = 2,  // PassThroughActivation<ap_uint<16>>         -> 16-bit scores (CNV last layer)
};

struct LayerDesc {
  int kind;
  int ifm_ch, ifm_dim, ofm_ch, ofm_dim;  // conv geometry; FC: dims 1, ifm_ch = K, ofm_ch = N
  int K, N;                              // MatrixW, MatrixH (convlayer.h:85-121: K = 9*IFM_CH)
  int pe, simd, wbits;                   // folding and weight precision (L*_PE, L*_SIMD, L*_WPI)
  int nthr;                              // thresholds per row the loader reads (0: CNV last layer)
  int api;                               // L*_API of config.h (fault topology only)
  int thr_bits;                          // 24 (CNV layer 0), 16, 0 (no threshold memory): topology activation_element_bits
  int in_mode, out_mode, pool;           // pool: 2x2 max-pool follows this layer (top.cpp:217,222)
  int wmem, tmem;                        // derived: (K/SIMD)*(N/PE), N/PE (mvau.hpp:100,104)
};

struct NetDesc {
  char name[16];
  int is_cnv;
  int nlayers;
  LayerDesc L[9];
  int in_bytes;    // per image, as BlackBoxJam reads it: CNV 3072 (int8 HWC), LFC 104 (832 bits)
  int out_bytes;   // CNV 128 (64 x int16), LFC 8
};

// Returns false for an unknown network name.
bool make_net(const char* name, NetDesc* out);

// Per-image element counts of the activation tensor a layer WRITES (after its pool), in 32-bit words.
inline int out_dim(const LayerDesc& l) { return l.pool ? l.ofm_dim / 2 : l.ofm_dim; }
inline int out_words_per_pixel(const LayerDesc& l) {
  if (l.out_mode == OUT_SCORE) return l.N / 2;             // int16 scores
  return (l.N / 32) * (l.out_mode == OUT_TERN ? 2 : 1);    // bit planes
}

}  // namespace bnnb
