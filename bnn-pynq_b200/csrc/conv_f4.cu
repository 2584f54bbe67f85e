// conv3x3 layers on tcgen05.mma kind::mxf4 (block-scaled fp4 with every scale = 1.0): an exact small-integer MMA at twice the
// int8 rate.  Activations {0,1} / {-1,0,+1} and weights {-1,+1} / {-2..+1} are all exactly representable in e2m1; products and
// the fp32 accumulation are exact (|acc| <= 2*2304 << 2^24), so the thresholded outputs are bit-identical to the reference's
// integer MVAU (mvau.hpp:86-172, activations.hpp:139-147).  Evidence: tests/test_gpu_tcgen05.py::test_fp4_probe_*.
//
// Same implicit GEMM as conv_tc.cu (operand planes of 16 bytes = 32 channels per pixel, tap (ky,kx) = descriptor start shifted by
// (ky*IDIM+kx) pixels), now one K=64 MMA per 64 input channels per tap, and a warp-specialised CTA (one per SM):
//   warps 0-3  epilogue: TMEM -> threshold -> bit rows in smem -> (pool) -> global
//   warps 4-7  expander: bit-packed activations of group g+1 (global, L2) -> fp4 operand planes, other buffer
//   warp  8    MMA issuer, up to three tiles ahead (three TMEM accumulator slots)
// decoupled by mbarriers only: abuf_full / abuf_empty per operand buffer, full / empty per accumulator slot.
#include <climits>
#include "kernels.cuh"
#include "tcgen05.cuh"

namespace bnnb {
namespace {

template <int C, int N, int IDIM, int OD, bool POOL, int G, int INM, int OUTM>
struct F4Cfg {
  static constexpr int PXI = IDIM * IDIM, ROWS = G * PXI, TILES = (ROWS + 127) / 128, MAXSHIFT = 2 * IDIM + 2;
  static constexpr int PITCH = (TILES * 128 + MAXSHIFT + 7) / 8 * 8;
  static constexpr int KC = C / 32;  // 16-byte chunks (32 fp4 channels) per tap
  static constexpr int PLANE = PITCH * 16, A_BYTES = KC * PLANE, B_BYTES = 9 * KC * N * 16;
  static constexpr int IPL = INM == IN_TERN ? 2 : 1, OPL = OUTM == OUT_TERN ? 2 : 1, NW = N / 32, WPP = NW * OPL;
  static constexpr int BITS_BYTES = TILES * 128 * WPP * 4;
  static constexpr int OFF_B = 2 * A_BYTES, OFF_BITS = OFF_B + B_BYTES, OFF_THR = OFF_BITS + BITS_BYTES, SMEM = OFF_THR + 2 * N * 4;
  static constexpr int SLOTS = 384 / N > 4 ? 4 : 384 / N;  // accumulator slots below the scale-factor columns
  static constexpr int SF_COL = 384, PD = POOL ? OD / 2 : OD;
};

__device__ __forceinline__ void hsync(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

// 8 bits -> 8 nibbles (bit i at position 4i)
__device__ __forceinline__ uint32_t spread8(uint32_t x) {
  x = (x | (x << 12)) & 0x000f000fu;
  x = (x | (x << 6)) & 0x03030303u;
  x = (x | (x << 3)) & 0x11111111u;
  return x;
}
// 32 channels (one word; ternary: nz word + pos word) -> 16 bytes of e2m1: 1 -> 0x2, -1 -> 0xA, 0 -> 0x0
template <int INM>
__device__ __forceinline__ uint4 expand_fp4(uint32_t w0, uint32_t w1) {
  uint32_t v[4];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const uint32_t a = spread8((w0 >> (8 * q)) & 0xff);
    if (INM == IN_TERN) {
      const uint32_t neg = spread8(((w0 & ~w1) >> (8 * q)) & 0xff);
      v[q] = (a << 1) | (neg << 3);
    } else {
      v[q] = a << 1;
    }
  }
  return make_uint4(v[0], v[1], v[2], v[3]);
}

template <int C, int N, int IDIM, int OD, bool POOL, int G, int INM, int OUTM>
__global__ void __launch_bounds__(288, 1) k_conv_f4(DevLayer L, const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int n_img,
                                                     int* __restrict__ status) {
  using Cfg = F4Cfg<C, N, IDIM, OD, POOL, G, INM, OUTM>;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[4], empty[4], abuf_full[2], abuf_empty[2];
  __shared__ uint32_t tmem_base;
  uint8_t* sb = smem + Cfg::OFF_B;
  uint32_t* sbits = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BITS);
  float* sthr = reinterpret_cast<float*>(smem + Cfg::OFF_THR);  // [2][N]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int role = warp < 4 ? 0 : (warp < 8 ? 1 : 2), ht = tid & 127, hw = warp & 3;
  constexpr int KCT = 9 * Cfg::KC;

  for (int i = tid; i < N * KCT; i += 288) {  // weights (packed e2m1, [N][K/2]) -> chunk-major operand layout
    const int n = i % N, kc = i / N;
    *reinterpret_cast<uint4*>(sb + ((size_t)kc * N + n) * 16) = *reinterpret_cast<const uint4*>(L.w4 + (size_t)n * (L.kpad / 2) + kc * 16);
  }
  for (int i = tid; i < 2 * N; i += 288) {
    const int t = L.thr_tc[i];
    sthr[(i & 1) * N + (i >> 1)] = t == INT_MAX ? 3.0e38f : (float)t;
  }
  for (int i = tid; i < 2 * Cfg::A_BYTES / 16; i += 288) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);  // pad pixels stay 0
  if (warp == 0) tc::tmem_alloc(&tmem_base, 512);
  if (tid == 0) {
    for (int i = 0; i < 2; i++) { tc::mbar_init(&abuf_full[i], 128); tc::mbar_init(&abuf_empty[i], 1); }
    for (int i = 0; i < 4; i++) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 4); }
    tc::mbar_init_fence();
  }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_base;
  if (warp < 4) tc::tmem_fill16(tmem + ((uint32_t)(warp * 32) << 16) + Cfg::SF_COL, 0x7f7f7f7fu);  // every scale factor = 1.0 (ue8m0)
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t idesc = tc::make_idesc_mxf4(128, N);
  const uint32_t sa_u0 = tc::smem_u32(smem), sb_u = tc::smem_u32(sb), sf = tmem + Cfg::SF_COL;
  bool failed = false;
  const int ngroups = (n_img + G - 1) / G;
  const int K = (int)blockIdx.x < ngroups ? (ngroups - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;  // groups of this CTA

  if (role == 1) {
    // ---- expander ------------------------------------------------------------------------------------------------------
    constexpr int CW = C / 32;
    for (int k = 0; k < K; k++) {
      const int img0 = ((int)blockIdx.x + k * (int)gridDim.x) * G, buf = k & 1;
      if (k >= 2 && !tc::mbar_wait(&abuf_empty[buf], (uint32_t)((k >> 1) - 1) & 1)) failed = true;
      uint8_t* sa = smem + (size_t)buf * Cfg::A_BYTES;
      for (int i = ht; i < Cfg::ROWS * CW; i += 128) {
        const int j = i % CW, m = i / CW, gi = m / Cfg::PXI;
        uint32_t w0 = 0, w1 = 0;
        if (img0 + gi < n_img) {
          const size_t base = ((size_t)(img0 + gi) * Cfg::PXI + (m - gi * Cfg::PXI)) * (CW * Cfg::IPL);
          w0 = __ldg(in + base + j);
          if (INM == IN_TERN) w1 = __ldg(in + base + CW + j);
        }
        *reinterpret_cast<uint4*>(sa + (size_t)j * Cfg::PLANE + (size_t)m * 16) = expand_fp4<INM>(w0, w1);
      }
      tc::fence_async_smem();
      tc::mbar_arrive(&abuf_full[buf]);
    }
  } else if (role == 2) {
    // ---- issuer ---------------------------------------------------------------------------------------------------------
    if (lane == 0) {
      for (int k = 0; k < K; k++) {
        const int buf = k & 1;
        if (!tc::mbar_wait(&abuf_full[buf], (uint32_t)(k >> 1) & 1)) failed = true;
        const uint32_t sa_u = sa_u0 + (uint32_t)buf * Cfg::A_BYTES;
        for (int t = 0; t < Cfg::TILES; t++) {
          const int gt = k * Cfg::TILES + t, slot = gt % Cfg::SLOTS, u = gt / Cfg::SLOTS;
          if (u >= 1 && !tc::mbar_wait(&empty[slot], (uint32_t)(u - 1) & 1)) failed = true;
          tc::fence_after_sync();
          const uint32_t acc = tmem + (uint32_t)slot * N;
          bool firstmma = true;
#pragma unroll
          for (int tap = 0; tap < 9; tap++) {
            const uint32_t off = (uint32_t)((tap / 3) * IDIM + tap % 3 + t * 128) * 16;
#pragma unroll
            for (int s = 0; s < Cfg::KC / 2; s++) {
              tc::mma_mxf4(acc, tc::make_desc(sa_u + (uint32_t)(2 * s) * Cfg::PLANE + off, Cfg::PLANE, 128),
                           tc::make_desc(sb_u + (uint32_t)((tap * Cfg::KC + 2 * s) * N) * 16, N * 16, 128), idesc, sf, sf, firstmma ? 0u : 1u);
              firstmma = false;
            }
          }
          tc::commit(&full[slot]);
        }
        tc::commit(&abuf_empty[buf]);
      }
    }
  } else {
    // ---- epilogue, pool, store ---------------------------------------------------------------------------------------------
    for (int k = 0; k < K; k++) {
      const int img0 = ((int)blockIdx.x + k * (int)gridDim.x) * G;
      for (int t = 0; t < Cfg::TILES; t++) {
        const int gt = k * Cfg::TILES + t, slot = gt % Cfg::SLOTS, u = gt / Cfg::SLOTS;
        if (!tc::mbar_wait(&full[slot], (uint32_t)u & 1)) failed = true;
        tc::fence_after_sync();
        const int m = t * 128 + hw * 32 + lane;
        const uint32_t trow = tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * N;
#pragma unroll
        for (int j = 0; j < Cfg::NW; j++) {
          uint32_t r[32];
          tc::tmem_ld32(trow + j * 32, r);
          tc::tmem_ld_wait();
          uint32_t b0 = 0, b1 = 0;
#pragma unroll
          for (int c = 0; c < 32; c++) {
            const float s = __uint_as_float(r[c]);
            const int t0 = sthr[j * 32 + c] < s;
            if (OUTM == OUT_BIT) b0 |= (uint32_t)t0 << c;
            else { const int t1 = sthr[N + j * 32 + c] < s; b0 |= (uint32_t)(1 ^ t0 ^ t1) << c; b1 |= (uint32_t)(t0 & t1) << c; }
          }
          sbits[(size_t)m * Cfg::WPP + j] = b0;
          if (OUTM == OUT_TERN) sbits[(size_t)m * Cfg::WPP + Cfg::NW + j] = b1;
        }
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&empty[slot]);
      }
      hsync(1);
      for (int i = ht; i < G * Cfg::PD * Cfg::PD * Cfg::NW; i += 128) {
        const int j = i % Cfg::NW, op = (i / Cfg::NW) % (Cfg::PD * Cfg::PD), gi = i / (Cfg::NW * Cfg::PD * Cfg::PD);
        if (img0 + gi >= n_img) continue;
        const int oy = op / Cfg::PD, ox = op % Cfg::PD;
        uint32_t o0, o1 = 0;
        if (POOL) {
          const uint32_t* r00 = sbits + (size_t)(gi * Cfg::PXI + (2 * oy) * IDIM + 2 * ox) * Cfg::WPP;
          const uint32_t *r01 = r00 + Cfg::WPP, *r10 = r00 + (size_t)IDIM * Cfg::WPP, *r11 = r10 + Cfg::WPP;
          if (OUTM == OUT_BIT) {
            o0 = r00[j] | r01[j] | r10[j] | r11[j];
          } else {
            const uint32_t P = r00[Cfg::NW + j] | r01[Cfg::NW + j] | r10[Cfg::NW + j] | r11[Cfg::NW + j];
            const uint32_t M = (r00[j] & ~r00[Cfg::NW + j]) & (r01[j] & ~r01[Cfg::NW + j]) & (r10[j] & ~r10[Cfg::NW + j]) & (r11[j] & ~r11[Cfg::NW + j]);
            o0 = P | M; o1 = P;
          }
        } else {
          const uint32_t* r = sbits + (size_t)(gi * Cfg::PXI + oy * IDIM + ox) * Cfg::WPP;
          o0 = r[j];
          if (OUTM == OUT_TERN) o1 = r[Cfg::NW + j];
        }
        uint32_t* dst = out + ((size_t)(img0 + gi) * (Cfg::PD * Cfg::PD) + op) * Cfg::WPP;
        dst[j] = o0;
        if (OUTM == OUT_TERN) dst[Cfg::NW + j] = o1;
      }
      hsync(1);
    }
  }
  if (failed && status) *status = 1;
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

template <int C, int N, int IDIM, int OD, bool POOL, int G, int INM, int OUTM>
bool launch_cfg(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int* status, cudaStream_t s) {
  using Cfg = F4Cfg<C, N, IDIM, OD, POOL, G, INM, OUTM>;
  static_assert(Cfg::SMEM <= 226 * 1024, "shared memory budget");
  static_assert(Cfg::SLOTS * N <= Cfg::SF_COL, "TMEM budget");
  auto kern = k_conv_f4<C, N, IDIM, OD, POOL, G, INM, OUTM>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
  const int ngroups = (n_img + G - 1) / G;
  kern<<<ngroups < num_sms ? ngroups : num_sms, 288, Cfg::SMEM, s>>>(L, in, out, n_img, status);
  return true;
}

}  // namespace

bool launch_conv_f4(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int* status, cudaStream_t s) {
  const LayerDesc& d = L.d;
  if (d.kind != KIND_CONV || !L.w4 || !L.thr_tc || n_img <= 0) return false;
  const bool a1 = d.in_mode == IN_XNOR && d.out_mode == OUT_BIT, a2 = d.in_mode == IN_TERN && d.out_mode == OUT_TERN;
  if (d.ifm_ch == 64 && d.N == 128 && d.ifm_dim == 14 && !d.pool) {
    if (a1) return launch_cfg<64, 128, 14, 12, false, 9, IN_XNOR, OUT_BIT>(L, in, out, n_img, num_sms, status, s);
    if (a2) return launch_cfg<64, 128, 14, 12, false, 6, IN_TERN, OUT_TERN>(L, in, out, n_img, num_sms, status, s);
  }
  if (d.ifm_ch == 128 && d.N == 128 && d.ifm_dim == 12 && d.pool) {
    if (a1) return launch_cfg<128, 128, 12, 10, true, 6, IN_XNOR, OUT_BIT>(L, in, out, n_img, num_sms, status, s);
    if (a2) return launch_cfg<128, 128, 12, 10, true, 5, IN_TERN, OUT_TERN>(L, in, out, n_img, num_sms, status, s);
  }
  return false;
}

}  // namespace bnnb
