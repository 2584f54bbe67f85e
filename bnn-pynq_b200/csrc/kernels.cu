// Hand-written sm_100a kernels: parameter re-pack, input preparation, the LOP3+POPC MVTU layers and class decode.
// Reference functions replaced (bnn/src/library/finn-hlslib): ConvLayer_Batch (convlayer.h:85-121),
// ConvolutionInputGenerator (slidingwindow.h:73-184), Matrix_Vector_Activate_Batch (mvau.hpp:86-172),
// mac/mul + XnorMul/Binary/Identity (mac.hpp:87-172, interpret.hpp:57-227), ThresholdsActivation (activations.hpp:126-148),
// StreamingMaxPool(_Precision)_Batch (maxpool.h:62-208), StreamingFCLayer_Batch (fclayer.h:83-111).
#include "kernels.cuh"
#include <climits>
#include <cstdio>

namespace bnnb {
thread_local bool tl_pdl = false;

// ------------------------------------------------------------------------------------------------------------
// Parameter re-pack: folded [pe][wmem] words -> K-major planes / int8 operands / int32 thresholds.
// row r = nf*PE + pe, column c = sf*SIMD + s, tile = nf*SF + sf (mvau.hpp:115,143-152,166-169).
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int weight_field(const uint64_t* __restrict__ wf, const LayerDesc& d, int r, int c) {
  const int pe = r % d.pe, nf = r / d.pe, sf = c / d.simd, s = c % d.simd;
  const uint64_t word = wf[(size_t)pe * d.wmem + nf * (d.K / d.simd) + sf];
  return (int)((word >> (s * d.wbits)) & ((1u << d.wbits) - 1));  // raw field: W1 bit, W2 two's-complement pair
}
__device__ __forceinline__ int weight_value(int field, int wbits) {
  if (wbits == 1) return field ? 1 : -1;   // Binary: 1 <-> +1, 0 <-> -1 (interpret.hpp:84,89)
  return (field ^ 2) - 2;                  // ap_int<2>: {0,1,2,3} -> {0,1,-2,-1} (weights.hpp:183-189)
}

__global__ void k_repack_planes(DevLayer L) {
  const LayerDesc& d = L.d;
  const int total = d.wbits * L.kwords * d.N;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int n = i % d.N, kw = (i / d.N) % L.kwords, plane = i / (d.N * L.kwords);
    uint32_t w = 0;
    for (int b = 0; b < 32; b++) w |= (uint32_t)((weight_field(L.wf, d, n, kw * 32 + b) >> plane) & 1) << b;
    L.wp[i] = w;
  }
}
// layer 0 of CNV: int8x4 words [64][8], K = 27 -> 32 (zero padded)
__global__ void k_repack_l0(DevLayer L) {
  const LayerDesc& d = L.d;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < d.N * 8; i += gridDim.x * blockDim.x) {
    const int n = i / 8, j = i % 8;
    uint32_t w = 0;
    for (int b = 0; b < 4; b++) {
      const int k = 4 * j + b;
      const int v = k < d.K ? weight_value(weight_field(L.wf, d, n, k), d.wbits) : 0;
      w |= (uint32_t)(v & 0xff) << (8 * b);
    }
    L.wp[i] = w;
  }
}
__global__ void k_repack_w8(DevLayer L) {
  const LayerDesc& d = L.d;
  const int total = d.N * L.kpad;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int n = i / L.kpad, k = i % L.kpad;
    L.w8[i] = k < d.K ? (int8_t)weight_value(weight_field(L.wf, d, n, k), d.wbits) : (int8_t)0;
  }
}
// e2m1 nibbles of the same values: 0 -> 0x0, +1 -> 0x2, -1 -> 0xA, -2 -> 0xC; even k in the low nibble
__global__ void k_repack_w4(DevLayer L) {
  const LayerDesc& d = L.d;
  const int total = d.N * (L.kpad / 2);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int n = i / (L.kpad / 2), kb = i % (L.kpad / 2);
    uint32_t byte = 0;
    for (int h = 0; h < 2; h++) {
      const int k = 2 * kb + h;
      const int v = k < d.K ? weight_value(weight_field(L.wf, d, n, k), d.wbits) : 0;
      const uint32_t nib = v == 0 ? 0x0u : (v == 1 ? 0x2u : (v == -1 ? 0xAu : 0xCu));
      byte |= nib << (4 * h);
    }
    L.w4[i] = (uint8_t)byte;
  }
}
// w4 -> slabs of NB channels x 64 K in the UMMA K-major operand layout (two 16-byte chunks per row), contiguous per (block, slab)
__global__ void k_repack_w4s(DevLayer L, int NB) {
  const LayerDesc& d = L.d;
  const int ks = L.kpad / 64, total = d.N * (L.kpad / 2) / 16;  // 16-byte units
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int n = i % NB, c = (i / NB) % 2, s = (i / (2 * NB)) % ks, nb = i / (2 * NB * ks);
    reinterpret_cast<uint4*>(L.w4s)[i] = *reinterpret_cast<const uint4*>(L.w4 + (size_t)(nb * NB + n) * (L.kpad / 2) + s * 32 + c * 16);
  }
}
__global__ void k_repack_thr(DevLayer L) {
  const LayerDesc& d = L.d;
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < d.N; n += gridDim.x * blockDim.x) {
    const int pe = n % d.pe, nf = n / d.pe;
    for (int k = 0; k < 2; k++)
      L.thr[2 * n + k] = k < d.nthr ? L.tf[((size_t)pe * d.tmem + nf) * d.nthr + k] : INT_MAX;
  }
}

void launch_repack(const DevLayer& L, int is_cnv_layer0, cudaStream_t s) {
  if (is_cnv_layer0) k_repack_l0<<<2, 256, 0, s>>>(L);
  else k_repack_planes<<<(L.d.wbits * L.kwords * L.d.N + 255) / 256, 256, 0, s>>>(L);
  if (L.w8) k_repack_w8<<<(L.d.N * L.kpad + 255) / 256, 256, 0, s>>>(L);
  if (L.w4) k_repack_w4<<<(L.d.N * (L.kpad / 2) + 255) / 256, 256, 0, s>>>(L);
  if (L.w4 && L.w4s && L.kpad % 64 == 0 && L.d.N % 64 == 0) k_repack_w4s<<<(L.d.N * (L.kpad / 32) + 255) / 256, 256, 0, s>>>(L, gemm_f4_block(L.d));
  k_repack_thr<<<(L.d.N + 255) / 256, 256, 0, s>>>(L);
  if (L.w8 && L.thr_tc) launch_thr_tc(L, s);
}

// ------------------------------------------------------------------------------------------------------------
// Fault epochs: one parameter word rewritten = the device operands derived from it rewritten in place (no re-pack of the layer).
// The patch carries the new folded word (the host mirror ran the reference's read-modify-write, foldedmv-offload.h:146-167); one
// warp applies the list IN ORDER (later entries for the same word win, like the reference's sequential RMWs):
//   weight word [pe][ind] = SIMD fields of row n = nf*PE + pe, columns [sf*SIMD, (sf+1)*SIMD), ind = nf*SF + sf (mvau.hpp:115,143-152)
//     -> popc planes (the <= 3 words per plane the columns fall into), int8 operand bytes, e2m1 operand bytes and their slab copy
//        (the 16-byte units holding them), layer-0 int8x4 words, and thr_tc[n] +- the change in the row's count of -1 weights
//   threshold [pe][ind][k] of row n = ind*PE + pe -> thr[n][k] and thr_tc[n][k] (same offset as before)
// ------------------------------------------------------------------------------------------------------------
__global__ void k_patch_apply(const DevLayer* __restrict__ Ls, int is_cnv, const DevPatch* __restrict__ p, int n) {
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  for (int i = 0; i < n; i++) {
    const DevPatch e = p[i];
    const DevLayer& L = Ls[e.layer];
    const LayerDesc& d = L.d;
    if (e.is_thr) {
      if (lane == 0 && d.nthr > 0) {
        const int nv = (int32_t)(int64_t)e.value;
        const_cast<int32_t*>(L.tf)[((size_t)e.pe * d.tmem + e.ind) * d.nthr + e.thresh] = nv;
        const int row = (int)e.ind * d.pe + (int)e.pe, k = (int)e.thresh;
        if (k < 2 && k < d.nthr) {
          const int old = L.thr[2 * row + k];
          L.thr[2 * row + k] = nv;
          if (L.thr_tc && d.out_mode != OUT_SCORE) L.thr_tc[2 * row + k] = nv - (old - L.thr_tc[2 * row + k]);  // thr_tc = thr - #neg
        }
      }
      __syncwarp();
      continue;
    }
    const int SF = d.K / d.simd, nf = (int)e.ind / SF, sf = (int)e.ind - nf * SF;
    const int row = nf * d.pe + (int)e.pe, k0 = sf * d.simd, k1 = k0 + d.simd;  // columns [k0, k1)
    uint64_t* wfw = const_cast<uint64_t*>(L.wf) + (size_t)e.pe * d.wmem + e.ind;
    const uint64_t oldw = *wfw;
    __syncwarp();
    if (lane == 0) *wfw = e.value;
    // Everything below derives from the patch's own value (the fields of columns k0..k1-1), never from the folded memory: a chain of
    // dependent global loads per bit made one patch cost tens of microseconds, and the fault path pays it once per fault.
    const uint32_t fmask = (1u << d.wbits) - 1;
    auto fld = [&](int k) { return (int)((e.value >> ((k - k0) * d.wbits)) & fmask); };  // k in [k0, k1)
    if (is_cnv && e.layer == 0) {  // int8x4 words [64][8], K = 27 padded to 32
      for (int k = k0 + lane; k < k1; k += 32) reinterpret_cast<int8_t*>(L.wp)[row * 32 + k] = (int8_t)weight_value(fld(k), d.wbits);
      __syncwarp();
      continue;
    }
    {  // popc planes: uint32 [plane][K/32][N]: the bits of the range replaced inside the (at most two) words it touches
      const int w0 = k0 >> 5, nwd = ((k1 - 1) >> 5) - w0 + 1;
      for (int j = lane; j < nwd * d.wbits; j += 32) {
        const int plane = j / nwd, kw = w0 + j % nwd;
        const int lo = max(k0, kw * 32), hi = min(k1, kw * 32 + 32);
        uint32_t bits = 0, m = 0;
        for (int k = lo; k < hi; k++) { bits |= (uint32_t)((fld(k) >> plane) & 1) << (k & 31); m |= 1u << (k & 31); }
        uint32_t* wp = L.wp + ((size_t)plane * L.kwords + kw) * d.N + row;
        *wp = (*wp & ~m) | bits;
      }
    }
    if (L.w8)
      for (int k = k0 + lane; k < k1; k += 32) L.w8[(size_t)row * L.kpad + k] = (int8_t)weight_value(fld(k), d.wbits);
    if (L.w4) {
      const int b0 = k0 >> 1, b1 = (k1 - 1) >> 1;  // bytes (two nibbles each); a nibble outside the range keeps its old code
      for (int kb = b0 + lane; kb <= b1; kb += 32) {
        uint8_t* bp = L.w4 + (size_t)row * (L.kpad / 2) + kb;
        uint32_t byte = *bp;
        for (int h = 0; h < 2; h++) {
          const int k = 2 * kb + h;
          if (k < k0 || k >= k1) continue;
          const int v = weight_value(fld(k), d.wbits);
          byte = (byte & ~(0xFu << (4 * h))) | ((v == 0 ? 0x0u : (v == 1 ? 0x2u : (v == -1 ? 0xAu : 0xCu))) << (4 * h));
        }
        *bp = (uint8_t)byte;
      }
      __syncwarp();
      if (L.w4s && L.kpad % 64 == 0 && d.N % 64 == 0) {  // slab copy: 16-byte units [N/NB][K/64][2][NB]
        const int NB = d.N % 128 == 0 ? 128 : 64, ks = L.kpad / 64;
        const int u0 = b0 >> 4, u1 = b1 >> 4;
        for (int u = u0 + lane; u <= u1; u += 32) {
          const int s = u >> 1, c = u & 1, nb = row / NB, nl = row - nb * NB;
          reinterpret_cast<uint4*>(L.w4s)[((size_t)(nb * ks + s) * 2 + c) * NB + nl] = *reinterpret_cast<const uint4*>(L.w4 + (size_t)row * (L.kpad / 2) + u * 16);
        }
      }
    }
    if (lane == 0 && L.thr_tc && d.in_mode == IN_XNOR) {  // 1-bit weights: #(-1) = zero bits of the field range
      const uint64_t mask = d.simd >= 64 ? ~0ull : ((1ull << d.simd) - 1);
      const int delta = __popcll(~e.value & mask) - __popcll(~oldw & mask);
      if (d.out_mode == OUT_SCORE) L.thr_tc[2 * row] += delta;
      else
        for (int k = 0; k < 2; k++)
          if (L.thr_tc[2 * row + k] != INT_MAX) L.thr_tc[2 * row + k] -= delta;
    }
    __syncwarp();
  }
}
void launch_patch_apply(const DevLayer* d_layers, int is_cnv, const DevPatch* patches, int n, cudaStream_t s) {
  if (n > 0) k_patch_apply<<<1, 32, 0, s>>>(d_layers, is_cnv, patches, n);
}

// ------------------------------------------------------------------------------------------------------------
// Input preparation
// ------------------------------------------------------------------------------------------------------------
// byte c -> f = -1 + 2c/255 (float) -> ap_fixed<8,1,AP_RND,AP_SAT> raw (cnvW1A1/sw/main_python.cpp:129, foldedmv-offload.h:139)
__device__ __forceinline__ int cifar_q(int c) {
  const float f = -1.0f + 2.0f * (float)c / 255.0f;
  float r = floorf(f * 128.0f + 0.5f);
  r = fminf(fmaxf(r, -128.0f), 127.0f);
  return (int)r;
}
// records [n][3073] (label, R plane, G plane, B plane) -> int8 [n][1024 px][3]   (chaninterleave + quantiseAndPack)
// One image per CTA iteration: the 3 073-byte record (not even 4-byte aligned) is staged in shared memory with aligned 32-bit loads,
// then every thread emits whole output words (4 interleaved bytes through the quantisation LUT), coalesced.
__global__ void __launch_bounds__(256) k_prep_cifar(const uint8_t* __restrict__ rec, int8_t* __restrict__ q, int n) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // the one-image graph: the first layer kernel may start its prologue
  __shared__ int8_t lut[256];
  __shared__ __align__(16) uint8_t buf[3073 + 7];
  lut[threadIdx.x & 255] = (int8_t)cifar_q(threadIdx.x & 255);
  for (int img = blockIdx.x; img < n; img += gridDim.x) {
    const uint8_t* src = rec + (size_t)img * 3073;
    const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 3);
    const uint32_t* src4 = reinterpret_cast<const uint32_t*>(src - mis);
    __syncthreads();  // previous image's reads of buf are done (and the LUT is written)
    for (int w = threadIdx.x; w < (3073 + 3 + 3) / 4; w += 256) reinterpret_cast<uint32_t*>(buf)[w] = __ldg(src4 + w);  // stays inside the 4-byte words the record touches
    __syncthreads();
    const uint8_t* p = buf + mis + 1;  // R plane; G at +1024, B at +2048
    uint32_t* dst = reinterpret_cast<uint32_t*>(q + (size_t)img * 3072);
    for (int w = threadIdx.x; w < 768; w += 256) {
      uint32_t v = 0;
#pragma unroll
      for (int e = 0; e < 4; e++) {
        const int o = 4 * w + e, px = o / 3, c = o - 3 * px;
        v |= (uint32_t)(uint8_t)lut[p[c * 1024 + px]] << (8 * e);
      }
      dst[w] = v;
    }
  }
}
void launch_prep_cifar(const uint8_t* records, int8_t* q, int n, cudaStream_t s) {
  if (n <= 0) return;
  k_prep_cifar<<<n < 148 * 8 ? n : 148 * 8, 256, 0, s>>>(records, q, n);
}
// pixels [n][784] -> bits [n][26 words]; bit = (byte >= 128)  <=>  (c/255)*2-1 >= 0 in float (foldedmv-offload.cpp:89-92)
__global__ void k_prep_mnist(const uint8_t* __restrict__ px, uint32_t* __restrict__ bits, int n) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // the one-image graph: the first layer kernel may start its prologue
  const size_t total = (size_t)n * 26;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const size_t img = i / 26;
    const int w = (int)(i % 26);
    uint32_t v = 0;
    for (int b = 0; b < 32; b++) {
      const int k = w * 32 + b;
      if (k < 784 && px[img * 784 + k] >= 128) v |= 1u << b;
    }
    bits[i] = v;
  }
}
void launch_prep_mnist(const uint8_t* pixels, uint32_t* bits, int n, cudaStream_t s) {
  if (n <= 0) return;
  const size_t total = (size_t)n * 26;
  const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  k_prep_mnist<<<blocks, 256, 0, s>>>(pixels, bits, n);
}

// ------------------------------------------------------------------------------------------------------------
// CNV layer 0: int8 pixels x {+-1 | 2-bit} weights, K = 27, accumulator = raw ap_fixed<24,16> = sum w * (2q)
// (cnvW1A1/hw/top.cpp:61,214).  One thread per output pixel, 64 channels looped with the weights broadcast from smem.
// ------------------------------------------------------------------------------------------------------------
template <int OUTM>
__global__ void __launch_bounds__(256) k_conv0(DevLayer L, const int8_t* __restrict__ in, uint32_t* __restrict__ out, int n_img,
                                                int32_t* __restrict__ tap) {
  __shared__ uint4 sw[64 * 2];
  __shared__ int sthr[128];
  for (int i = threadIdx.x; i < 64 * 8; i += blockDim.x) ((uint32_t*)sw)[i] = L.wp[i];
  for (int i = threadIdx.x; i < 128; i += blockDim.x) sthr[i] = L.thr[i];
  __syncthreads();
  const long long total = (long long)n_img * 900;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (long long)gridDim.x * blockDim.x) {
    const long long img = p / 900;
    const int pix = (int)(p % 900), oy = pix / 30, ox = pix % 30;
    const int8_t* src = in + img * 3072 + (oy * 32 + ox) * 3;
    // gather the 27 window bytes, column order (ky*3+kx)*3 + c (slidingwindow.h:133-155), packed 4 per word
    uint32_t x[7];
#pragma unroll
    for (int j = 0; j < 7; j++) x[j] = 0;
#pragma unroll
    for (int k = 0; k < 27; k++) {
      const int ky = k / 9, r = k % 9;
      const uint32_t b = (uint8_t)__ldg(src + ky * 96 + r);
      x[k >> 2] |= b << (8 * (k & 3));
    }
    uint32_t b0[2] = {0, 0}, b1[2] = {0, 0};
#pragma unroll 4
    for (int n = 0; n < 64; n++) {
      const uint4 wa = sw[2 * n], wb = sw[2 * n + 1];
      int d = __dp4a((int)x[0], (int)wa.x, 0);
      d = __dp4a((int)x[1], (int)wa.y, d);
      d = __dp4a((int)x[2], (int)wa.z, d);
      d = __dp4a((int)x[3], (int)wa.w, d);
      d = __dp4a((int)x[4], (int)wb.x, d);
      d = __dp4a((int)x[5], (int)wb.y, d);
      d = __dp4a((int)x[6], (int)wb.z, d);
      const int acc = 2 * d;
      if (tap) tap[p * 64 + n] = acc;
      const int t0 = sthr[2 * n] < acc;  // strict compare (activations.hpp:141-146)
      if (OUTM == OUT_BIT) {
        b0[n >> 5] |= (uint32_t)t0 << (n & 31);
      } else {
        const int t1 = sthr[2 * n + 1] < acc;
        b0[n >> 5] |= (uint32_t)(1 ^ t0 ^ t1) << (n & 31);  // nz: value = -1 + t0 + t1 != 0
        b1[n >> 5] |= (uint32_t)(t0 & t1) << (n & 31);      // pos
      }
    }
    if (OUTM == OUT_BIT) {
      reinterpret_cast<uint2*>(out)[p] = make_uint2(b0[0], b0[1]);
    } else {
      reinterpret_cast<uint4*>(out)[p] = make_uint4(b0[0], b0[1], b1[0], b1[1]);
    }
  }
}
void launch_conv0(const DevLayer& L, const int8_t* in, uint32_t* out, int n_img, int32_t* tap, int num_sms, cudaStream_t s) {
  if (n_img <= 0) return;
  long long total = (long long)n_img * 900;
  int blocks = (int)((total + 255) / 256 < (long long)num_sms * 8 ? (total + 255) / 256 : (long long)num_sms * 8);
  if (L.d.out_mode == OUT_BIT) k_conv0<OUT_BIT><<<blocks, 256, 0, s>>>(L, in, out, n_img, tap);
  else k_conv0<OUT_TERN><<<blocks, 256, 0, s>>>(L, in, out, n_img, tap);
}

// ------------------------------------------------------------------------------------------------------------
// Generic MVTU layer on CUDA cores (LOP3 + POPC): conv3x3 "valid" / FC, optional fused 2x2 max-pool,
// 1-bit or ternary inputs, 1- or 2-bit weights.  Persistent CTAs keep a 32*NCL-channel slice of the layer's weight planes
// in shared memory; a warp owns a group of 4 rows (4 conv pixels of one pooled pixel, or 4 consecutive rows), lanes own
// channels, and threshold bits are packed with warp ballots.
//   IN_XNOR          acc = sum popc(~(x ^ w))                                   (interpret.hpp:57-69)
//   IN_BIN           acc = 2*sum popc(~(x ^ w)) - K                             (interpret.hpp:100-103)
//   IN_TERN, W1      acc = 2*sum popc(nz & ~(w ^ pos)) - sum popc(nz)           (interpret.hpp:86-90)
//   IN_TERN, W2      acc = sum [popc(l&P) - popc(l&M)] - 2*sum [popc(s&P) - popc(s&M)],  P = nz&pos, M = nz&~pos
// ------------------------------------------------------------------------------------------------------------
template <int NCL, int WB, int INM>
__global__ void __launch_bounds__(256) k_layer_popc(DevLayer L, const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int n_img,
                                                     int32_t* __restrict__ tap) {
  extern __shared__ uint32_t sw[];  // [WB][kwords][NCL][32]
  const LayerDesc& d = L.d;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int chunk = blockIdx.y, cbase = chunk * 32 * NCL;
  const int kwords = L.kwords;
  for (int i = threadIdx.x; i < WB * kwords * NCL * 32; i += blockDim.x) {
    const int l = i & 31, j = (i >> 5) % NCL, kw = (i / (32 * NCL)) % kwords, p = i / (32 * NCL * kwords);
    sw[i] = L.wp[((size_t)p * kwords + kw) * d.N + cbase + j * 32 + l];
  }
  int thr0[NCL], thr1[NCL];
#pragma unroll
  for (int j = 0; j < NCL; j++) {
    thr0[j] = L.thr[2 * (cbase + j * 32 + lane)];
    thr1[j] = L.thr[2 * (cbase + j * 32 + lane) + 1];
  }
  __syncthreads();

  constexpr int PL = (INM == IN_TERN) ? 2 : 1;            // input planes
  const int cw = d.kind == KIND_CONV ? d.ifm_ch / 32 : kwords;  // words per pixel per plane
  const int taps = d.kind == KIND_CONV ? 9 : 1;
  const int id = d.ifm_dim, od = d.ofm_dim, opix = od * od;
  const int in_img_words = id * id * cw * PL;
  const int pd = d.pool ? od / 2 : od;
  const int nw = d.N / 32;
  const int out_planes = d.out_mode == OUT_TERN ? 2 : 1;
  const long long rows = (long long)n_img * opix;
  const long long groups = d.pool ? (long long)n_img * pd * pd : (rows + 3) / 4;

  for (long long g = (long long)blockIdx.x * nwarps + warp; g < groups; g += (long long)gridDim.x * nwarps) {
    long long rimg[4];
    int rpix[4];
    bool valid[4];
    long long base[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
      long long r;
      if (d.pool) {
        const long long img = g / (pd * pd);
        const int pp = (int)(g % (pd * pd)), py = pp / pd, px = pp % pd;
        rimg[i] = img; rpix[i] = (2 * py + (i >> 1)) * od + 2 * px + (i & 1); valid[i] = true;
        r = img * opix + rpix[i];
      } else {
        r = g * 4 + i;
        valid[i] = r < rows;
        if (!valid[i]) r = rows - 1;
        rimg[i] = r / opix; rpix[i] = (int)(r % opix);
      }
      const int oy = rpix[i] / od, ox = rpix[i] % od;
      base[i] = rimg[i] * in_img_words + (long long)(oy * id + ox) * cw * PL;
    }
    int acc[4][NCL];
    int nzc[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < NCL; j++) acc[i][j] = 0;

    for (int t = 0; t < taps; t++) {
      const int toff = ((t / 3) * id + (t % 3)) * cw * PL;
      for (int c = 0; c < cw; c++) {
        const int kw = t * cw + c;
        uint32_t x[4], xp[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
          x[i] = __ldg(in + base[i] + toff + c);
          if (PL == 2) xp[i] = __ldg(in + base[i] + toff + cw + c);
        }
        if (INM == IN_TERN) {
#pragma unroll
          for (int i = 0; i < 4; i++) {
            if (WB == 1) nzc[i] += __popc(x[i]);
            else { const uint32_t nz = x[i]; x[i] = nz & xp[i]; xp[i] = nz & ~xp[i]; }  // P, M
          }
        }
#pragma unroll
        for (int j = 0; j < NCL; j++) {
          const uint32_t w0 = sw[(kw * NCL + j) * 32 + lane];
          uint32_t w1 = 0;
          if (WB == 2) w1 = sw[((kwords + kw) * NCL + j) * 32 + lane];
#pragma unroll
          for (int i = 0; i < 4; i++) {
            if (INM == IN_XNOR || INM == IN_BIN) acc[i][j] += __popc(~(x[i] ^ w0));
            else if (WB == 1) acc[i][j] += __popc(x[i] & ~(w0 ^ xp[i]));
            else acc[i][j] += __popc(w0 & x[i]) - __popc(w0 & xp[i]) - 2 * (__popc(w1 & x[i]) - __popc(w1 & xp[i]));
          }
        }
      }
    }
    // finish the accumulators
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = 0; j < NCL; j++) {
        if (INM == IN_BIN) acc[i][j] = 2 * acc[i][j] - d.K;
        if (INM == IN_TERN && WB == 1) acc[i][j] = 2 * acc[i][j] - nzc[i];
      }
    if (tap) {
#pragma unroll
      for (int i = 0; i < 4; i++)
        if (valid[i])
#pragma unroll
          for (int j = 0; j < NCL; j++) tap[(rimg[i] * opix + rpix[i]) * d.N + cbase + j * 32 + lane] = acc[i][j];
    }
    if (d.out_mode == OUT_SCORE) {  // PassThroughActivation<ap_uint<16>> (activations.hpp:85-92): acc mod 2^16
      int16_t* o16 = reinterpret_cast<int16_t*>(out);
#pragma unroll
      for (int i = 0; i < 4; i++)
        if (valid[i])
#pragma unroll
          for (int j = 0; j < NCL; j++) o16[(rimg[i] * opix + rpix[i]) * d.N + cbase + j * 32 + lane] = (int16_t)(acc[i][j] & 0xffff);
      continue;
    }
    // thresholds (strict <, on the 16-bit accumulator) -> bits -> pool -> ballot
#pragma unroll
    for (int j = 0; j < NCL; j++) {
      uint32_t pos[4], neg[4];
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int a = (int)(int16_t)acc[i][j];
        const int t0 = thr0[j] < a, t1 = thr1[j] < a;
        if (d.out_mode == OUT_BIT) { pos[i] = t0; neg[i] = 0; }
        else { pos[i] = t0 & t1; neg[i] = 1 ^ (t0 | t1); }
      }
      if (d.pool) {
        // A1: OR (maxpool.h:77-83); ternary: signed max with init -1 (maxpool.h:141,156-160)
        const uint32_t P = pos[0] | pos[1] | pos[2] | pos[3], M = neg[0] & neg[1] & neg[2] & neg[3];
        const long long o = (rimg[0] * (pd * pd) + (rpix[0] / od / 2) * pd + (rpix[0] % od) / 2) * nw * out_planes;
        if (d.out_mode == OUT_BIT) {
          const uint32_t wv = __ballot_sync(0xffffffffu, P);
          if (lane == 0) out[o + chunk * NCL + j] = wv;
        } else {
          const uint32_t wnz = __ballot_sync(0xffffffffu, P | M), wpos = __ballot_sync(0xffffffffu, P);
          if (lane == 0) { out[o + chunk * NCL + j] = wnz; out[o + nw + chunk * NCL + j] = wpos; }
        }
      } else {
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const long long o = (rimg[i] * opix + rpix[i]) * nw * out_planes;
          if (d.out_mode == OUT_BIT) {
            const uint32_t wv = __ballot_sync(0xffffffffu, pos[i]);
            if (lane == 0 && valid[i]) out[o + chunk * NCL + j] = wv;
          } else {
            const uint32_t wnz = __ballot_sync(0xffffffffu, pos[i] | neg[i]), wpos = __ballot_sync(0xffffffffu, pos[i]);
            if (lane == 0 && valid[i]) { out[o + chunk * NCL + j] = wnz; out[o + nw + chunk * NCL + j] = wpos; }
          }
        }
      }
    }
  }
}

template <int NCL, int WB, int INM>
static void launch_popc_t(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int32_t* tap, int num_sms, cudaStream_t s) {
  const size_t smem = (size_t)WB * L.kwords * NCL * 32 * 4;
  static bool attr_done = false;  // per instantiation; a process uses one device per engine instance set (attribute is per device)
  (void)attr_done;
  cudaFuncSetAttribute(k_layer_popc<NCL, WB, INM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  const int chunks = L.d.N / (32 * NCL);
  int occ = (int)((220 * 1024) / (smem + 1024));
  if (occ > 4) occ = 4;
  if (occ < 1) occ = 1;
  int gx = (num_sms * occ) / chunks;
  if (gx < 1) gx = 1;
  const long long rows = (long long)n_img * L.d.ofm_dim * L.d.ofm_dim;
  const long long groups = L.d.pool ? rows / 4 : (rows + 3) / 4;
  const long long need = (groups + 7) / 8;
  if (gx > need) gx = (int)(need < 1 ? 1 : need);
  k_layer_popc<NCL, WB, INM><<<dim3(gx, chunks), 256, smem, s>>>(L, in, out, n_img, tap);
}

void launch_layer_popc(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int32_t* tap, int num_sms, cudaStream_t s) {
  if (n_img <= 0) return;
  const int N = L.d.N, wb = L.d.wbits, im = L.d.in_mode;
#define BNNB_DISPATCH(NCL)                                                                                         \
  do {                                                                                                             \
    if (im == IN_XNOR) launch_popc_t<NCL, 1, IN_XNOR>(L, in, out, n_img, tap, num_sms, s);                          \
    else if (im == IN_BIN) launch_popc_t<NCL, 1, IN_BIN>(L, in, out, n_img, tap, num_sms, s);                       \
    else if (wb == 1) launch_popc_t<NCL, 1, IN_TERN>(L, in, out, n_img, tap, num_sms, s);                           \
    else launch_popc_t<NCL, 2, IN_TERN>(L, in, out, n_img, tap, num_sms, s);                                        \
  } while (0)
  if (N == 64) BNNB_DISPATCH(2);
  else if (N == 128) BNNB_DISPATCH(4);
  else BNNB_DISPATCH(8);  // 256, 512, 1024: 256-channel slices
#undef BNNB_DISPATCH
}

// ------------------------------------------------------------------------------------------------------------
// CNV class decode, multi-image rule: first score strictly above a running maximum that starts at 0
// (testPrebuiltCIFAR10_multiple_images, foldedmv-offload.h:398-406).
// ------------------------------------------------------------------------------------------------------------
__global__ void k_cnv_classes(const int16_t* __restrict__ sc, int n, int ncls, int* __restrict__ cls) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    int mi = 0, mv = 0;
    for (int j = 0; j < ncls; j++) {
      const int v = sc[(size_t)i * 64 + j];
      if (v > mv) { mv = v; mi = j; }
    }
    cls[i] = mi;
  }
}
void launch_cnv_classes(const int16_t* scores, int n, int ncls, int* classes, cudaStream_t s) {
  if (n <= 0) return;
  ncls = ncls < 0 ? 0 : (ncls > 64 ? 64 : ncls);  // BlackBoxJam writes 64 scores per image (cnvW1A1/hw/top.cpp:197)
  k_cnv_classes<<<(n + 255) / 256, 256, 0, s>>>(scores, n, ncls, classes);
}

// LFC class decode on the device, multi-image rule (foldedmv-offload.cpp:202-220): (unsigned) log2(word & mask), 0 when no bit is set;
// (below 2^32 the index of the top bit; above, floor(log2) of the word rounded to double, as the reference's libm call gives it)
__global__ void k_lfc_classes(const uint64_t* __restrict__ words, int n, uint64_t mask, int* __restrict__ cls) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint64_t w = words[i] & mask;
    int c = 0;
    if (w) c = w < (1ull << 32) ? 31 - __clz((unsigned)w) : (int)(unsigned)log2((double)w);
    cls[i] = c;
  }
}
void launch_lfc_classes(const uint64_t* words, int n, int ncls, int* classes, cudaStream_t s) {
  if (n <= 0) return;
  const uint64_t mask = ncls >= 64 ? ~0ull : (ncls <= 0 ? 0 : (~0ull >> (64 - ncls)));
  k_lfc_classes<<<(n + 255) / 256, 256, 0, s>>>(words, n, mask, classes);
}

void kernels_init_device() {}

}  // namespace bnnb
