// Device-side data layout and kernel launch wrappers (sm_100a).  See DESIGN.md "Data layout in HBM".
//
//  activations, 1 bit   : uint32 [img][pixel][C/32]            bit b of word j = channel 32j+b (1 <-> +1)
//  activations, ternary : uint32 [img][pixel][2][C/32]         plane 0 = nz (value != 0), plane 1 = pos (value == +1)
//  CNV input            : int8   [img][32][32][3]              quantiseAndPack<8,1> output (foldedmv-offload.h:129-144)
//  LFC input            : uint32 [img][26]                     binarizeAndPack output, 832 bits (foldedmv-offload.cpp:82-98)
//  CNV output           : int16  [img][64]                     PassThroughActivation scores (what Stream2Mem_Batch<64,128> writes)
//  LFC output           : uint64 [img]
//  weights (popc path)  : uint32 [plane][K/32][N]              K-major bit planes; W1: plane 0 = bit(+1); W2: plane 0 = lsb, 1 = sign
//  weights layer 0      : uint32 [64][8]                       int8x4 words, K = 27 padded to 32, values +-1 (W1) or -2..1 (W2)
//  weights (mma path)   : int8   [N][Kpad]                     values, K contiguous ("col" B operand of mma.m16n8k32)
//  thresholds           : int32  [N][2]                        sign-extended int16 (layer 0: raw ap_fixed<24,16>); unused slots INT_MAX
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "netdesc.hpp"

namespace bnnb {

struct DevLayer {
  LayerDesc d;
  const uint64_t* wf;   // folded weight memory  [pe][wmem]   (device copy of the reference-coordinates mirror)
  const int32_t* tf;    // folded threshold memory [pe][tmem][nthr]
  uint32_t* wp;         // popc-path planes (or layer-0 int8x4 words)
  int8_t* w8;           // mma-path int8 B operand (may be null)
  uint8_t* w4;          // e2m1 (fp4) B operand, two per byte, [N][kpad/2] (may be null)
  uint8_t* w4s;         // the same, as K=64 slabs in UMMA operand layout [N/NB][K/64][2][NB][16 B] for gemm_f4 (may be null)
  int32_t* thr;         // [N][2]
  int32_t* thr_tc;      // [N][2] thresholds adjusted for the tensor-core path (may be null)
  int kwords;           // K/32
  int kpad;             // K padded to a multiple of 32 (mma path)
};

// ---- launches -------------------------------------------------------------------------------------------------------------------
// The one-image CUDA graph (engine.cu run_single) launches the layer kernels with programmatic dependent launch: every kernel of
// the chain is resident at once, runs its prologue (weights and thresholds into shared memory, TMEM, barriers) and waits at
// griddepcontrol.wait, in the thread that first reads the layer below, for that layer to finish.  Set per host thread around
// run_chunk; every other path launches ordinarily (whole-GPU grids gain nothing, and early CTAs get in other lanes' way).
extern thread_local bool tl_pdl;
#ifdef __CUDACC__
template <typename... P, typename... A>
inline void launch_maybe_pdl(void (*kern)(P...), int grid, int block, size_t smem, cudaStream_t s, A... args) {
  if (!tl_pdl) {
    kern<<<grid, block, smem, s>>>(args...);
    return;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  cudaLaunchKernelEx(&cfg, kern, P(args)...);
}
#endif

// ---- parameter re-pack (runs once per load_parameters and after every fault epoch for the touched layers) ----
void launch_repack(const DevLayer& L, int is_cnv_layer0, cudaStream_t s);
// one rewritten parameter word: {layer,is_thr,pe,ind,thresh,pad,value}
struct DevPatch { uint32_t layer, is_thr, pe, ind, thresh, pad; uint64_t value; };
// applies the list in order to the folded mirrors AND to every operand derived from the touched words (no re-pack of the layer);
// d_layers: device copy of the network's DevLayer table
void launch_patch_apply(const DevLayer* d_layers, int is_cnv, const DevPatch* patches, int n, cudaStream_t s);

// ---- input preparation ----
void launch_prep_cifar(const uint8_t* records, int8_t* q, int n, cudaStream_t s);   // 3073-byte records -> int8 HWC
void launch_prep_mnist(const uint8_t* pixels, uint32_t* bits, int n, cudaStream_t s);  // 784 bytes -> 832 bits

// ---- layers (LOP3+POPC CUDA-core path) ----
// tap_acc (optional): int32 [img][ofm_dim^2][N] pre-threshold accumulators.
void launch_conv0(const DevLayer& L, const int8_t* in, uint32_t* out, int n_img, int32_t* tap_acc, int num_sms, cudaStream_t s);
void launch_layer_popc(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int32_t* tap_acc, int num_sms, cudaStream_t s);

// ---- layers (tcgen05 kind::i8 tensor-core path, conv_tc.cu); returns false when the layer shape has no instantiation ----
bool launch_conv_tc(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int* status, int32_t* tap, cudaStream_t s);
void launch_thr_tc(const DevLayer& L, cudaStream_t s);
// conv layers on kind::mxf4 (conv_f4.cu)
// in_fmt / out_fmt: ActFmt of f4_common.cuh (ACT_BITS everywhere except between fp4 kernels of a 1-bit network)
bool conv_f4_supports(const LayerDesc& d);
// tap (parity tests; null in production): int32 MMA sums [img][ofm_dim^2][N], pre-loads un-folded
// og: images per group of the consumer when out_fmt == ACT_F4I (conv_f4_group of the next layer)
bool launch_conv_f4(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int in_fmt, int out_fmt, int og, int* status, int32_t* tap,
                    cudaStream_t s);
int conv_f4_group(const LayerDesc& d);
// fully-connected layers on kind::mxf4 (gemm_f4.cu); gemm_f4_block = channel block width NB of the slab layout
int gemm_f4_block(const LayerDesc& d);
bool gemm_f4_supports(const DevLayer& L);
// classes / ncls: a score layer (CNV layer 8) also writes the class of every image (the multi-image arg-max rule), saving the launch of
// k_cnv_classes
bool launch_gemm_f4(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int in_fmt, int out_fmt, int* status, int32_t* tap,
                    cudaStream_t s, int* classes = nullptr, int ncls = 0);
// a run of gemm_f4 layers ending in the network's output layer, in one launch (fc_chain.cu): L[0..nl), first operand as ACT_F4T tiles;
// false when the run does not qualify (the caller launches the layers one by one)
bool launch_fc_chain(const DevLayer* L, int nl, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int in_fmt, int* status, cudaStream_t s,
                     int* classes, int ncls);
// K-streamed GEMM for FC / small conv layers (gemm_tc.cu)
bool launch_gemm_tc(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int* status, int32_t* tap, cudaStream_t s);
// fused CNV layers 0+1 (cnv01_tc.cu): int8 HWC image -> pooled, bit-packed layer-1 output
// Tap01 (parity tests; all null in production): acc0 / acc1 = int32 MMA sums [img][900][64] / [img][784][64] (layer 0: sum w*q;
// layer 1: sum a*w over the operand values), act0 = the layer-1 operand rows layer 0 wrote, uint4 [img][900][4]
// (fp4: entries 0 / 2 = channels 0-31 / 32-63 as e2m1 codes, FOLD nibble order; int8: four 16-channel chunks).
struct Tap01 { int32_t* acc0 = nullptr; int32_t* acc1 = nullptr; uint4* act0 = nullptr; };
bool launch_cnv01_tc(const DevLayer& L0, const DevLayer& L1, const int8_t* in, uint32_t* out, int n_img, int num_sms, bool fp4, int out_fmt, int* status,
                     const Tap01& tp, cudaStream_t s);

// ---- class decode for CNV (multi-image rule, foldedmv-offload.h:398-406) ----
void launch_cnv_classes(const int16_t* scores, int n, int ncls, int* classes, cudaStream_t s);
void launch_lfc_classes(const uint64_t* words, int n, int ncls, int* classes, cudaStream_t s);

// opt-in to large dynamic shared memory for every kernel (call once per device)
void kernels_init_device();

}  // namespace bnnb
