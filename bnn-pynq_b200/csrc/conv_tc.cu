// Tensor-core MVTU: conv3x3 "valid" layers as implicit GEMMs on tcgen05.mma kind::i8 (sm_100a), accumulators in TMEM.
//
// Replaces, for the layers it is selected for, ConvolutionInputGenerator + Matrix_Vector_Activate_Batch + ThresholdsActivation
// (+ StreamingMaxPool_Batch) of finn-hlslib (slidingwindow.h:73-184, mvau.hpp:86-172, activations.hpp:126-148, maxpool.h:62-208).
//
// Per CTA (persistent over groups of G images):
//   1. expand the group's bit-packed activations (global, L2-resident) to int8 in shared memory, laid out as the K-major,
//      un-swizzled UMMA operand: 16-channel chunk planes, pixel-major inside a plane (16 bytes per pixel per plane);
//   2. the sliding window is never materialised: for tap (ky,kx) the A operand of a 128-row tile is the SAME plane read from
//      a start address shifted by (ky*IDIM + kx) pixels (16 bytes each).  Rows are "full-width" pixel indices over the
//      concatenated images; rows whose window would wrap (x >= OD or y >= OD) are computed and discarded;
//   3. one elected thread issues 9*C/32 MMAs (M=128, N, K=32) per tile into one of two TMEM accumulators and commits to an
//      mbarrier; the four warps run the epilogue of tile t while the tensor core works on tile t+1;
//   4. epilogue: tcgen05.ld (one row per thread) -> integer threshold compare -> bits packed in registers (no ballots)
//      -> shared-memory bit rows -> 2x2 max-pool (OR / ternary max) -> global, bit-packed.
// The weights (int8 {-1,+1} or {-2..1}, [N][K]) stay in shared memory for the CTA's lifetime.
// Arithmetic: A1 inputs are expanded to {0,1}: S = sum a*w and the reference accumulator (#xnor matches) is S + #(w == -1),
// so the compare  thr < acc  becomes  thr - #neg(n) < S  (thresholds pre-adjusted on device by k_thr_tc).  Ternary inputs are
// expanded to {-1,0,+1}: S is the reference accumulator itself.
#include <climits>
#include "kernels.cuh"
#include "tcgen05.cuh"

namespace bnnb {

// adjusted thresholds for the tensor-core path: thrT[n][k] = thr[n][k] - (#negative weights of row n) for 1-bit inputs
__global__ void k_thr_tc(DevLayer L) {
  const LayerDesc& d = L.d;
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < d.N; n += gridDim.x * blockDim.x) {
    int neg = 0;
    if (d.in_mode == IN_XNOR)
      for (int k = 0; k < d.K; k++) neg += L.w8[(size_t)n * L.kpad + k] < 0;
    if (d.out_mode == OUT_SCORE) {  // no threshold: the epilogue adds this offset to S to get the reference accumulator
      L.thr_tc[2 * n] = neg;
      L.thr_tc[2 * n + 1] = 0;
      continue;
    }
    for (int k = 0; k < 2; k++) {
      const int t = L.thr[2 * n + k];
      L.thr_tc[2 * n + k] = t == INT_MAX ? INT_MAX : t - neg;
    }
  }
}
void launch_thr_tc(const DevLayer& L, cudaStream_t s) { k_thr_tc<<<(L.d.N + 127) / 128, 128, 0, s>>>(L); }

template <int C, int N, int IDIM, int OD, bool POOL, int G, int INM, int OUTM>
struct ConvTcCfg {
  static constexpr int PXI = IDIM * IDIM;
  static constexpr int ROWS = G * PXI;
  static constexpr int TILES = (ROWS + 127) / 128;
  static constexpr int MAXSHIFT = 2 * IDIM + 2;
  static constexpr int PITCH = (TILES * 128 + MAXSHIFT + 7) / 8 * 8;  // pixels per plane
  static constexpr int KC = C / 16;                                   // 16-byte chunks per tap
  static constexpr int PLANE = PITCH * 16;
  static constexpr int A_BYTES = KC * PLANE;
  static constexpr int B_BYTES = 9 * KC * N * 16;
  static constexpr int IPL = INM == IN_TERN ? 2 : 1;                  // input planes
  static constexpr int OPL = OUTM == OUT_TERN ? 2 : 1;
  static constexpr int NW = N / 32;
  static constexpr int BITS_BYTES = TILES * 128 * NW * OPL * 4;
  static constexpr int THR_BYTES = N * 2 * 4;
  static constexpr int SMEM = A_BYTES + B_BYTES + BITS_BYTES + THR_BYTES;
  static constexpr int TMEM_COLS = 2 * N < 32 ? 32 : 2 * N;
  static constexpr int PD = POOL ? OD / 2 : OD;
};

__device__ __forceinline__ uint32_t spread4(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }  // 4 bits -> 4 bytes of 0/1

template <int C, int N, int IDIM, int OD, bool POOL, int G, int INM, int OUTM>
__global__ void __launch_bounds__(128) k_conv_tc(DevLayer L, const uint32_t* __restrict__ in,
                                                  uint32_t* __restrict__ out, int n_img, int* __restrict__ status, int32_t* __restrict__ tap) {
  using Cfg = ConvTcCfg<C, N, IDIM, OD, POOL, G, INM, OUTM>;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ uint32_t tmem_base;
  uint8_t* sa = smem;
  uint8_t* sb = smem + Cfg::A_BYTES;
  uint32_t* sbits = reinterpret_cast<uint32_t*>(smem + Cfg::A_BYTES + Cfg::B_BYTES);
  int32_t* sthr = reinterpret_cast<int32_t*>(smem + Cfg::A_BYTES + Cfg::B_BYTES + Cfg::BITS_BYTES);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  constexpr int KCT = 9 * Cfg::KC;  // chunks over the whole K

  // ---- one-time setup: weights -> smem operand layout, thresholds, TMEM, barriers -------------------------------------
  for (int i = tid; i < N * KCT; i += 128) {
    const int n = i % N, kc = i / N;
    *reinterpret_cast<uint4*>(sb + ((size_t)kc * N + n) * 16) = *reinterpret_cast<const uint4*>(L.w8 + (size_t)n * L.kpad + kc * 16);
  }
  for (int i = tid; i < 2 * N; i += 128) sthr[(i & 1) * N + (i >> 1)] = L.thr_tc[i];  // [2][N]: consecutive channels -> LDS.128
  // the pad pixels of every plane are read by the last tile's shifted windows: keep them defined
  for (int i = tid; i < Cfg::KC * (Cfg::PITCH - Cfg::ROWS); i += 128) {
    const int g = i / (Cfg::PITCH - Cfg::ROWS), p = Cfg::ROWS + i % (Cfg::PITCH - Cfg::ROWS);
    *reinterpret_cast<uint4*>(sa + (size_t)g * Cfg::PLANE + p * 16) = make_uint4(0, 0, 0, 0);
  }
  if (warp == 0) tc::tmem_alloc(&tmem_base, Cfg::TMEM_COLS);
  if (tid == 0) { tc::mbar_init(&bar[0], 1); tc::mbar_init(&bar[1], 1); tc::mbar_init_fence(); }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_base;
  const uint32_t idesc = tc::make_idesc_i8(128, N);
  const uint32_t sa_u = tc::smem_u32(sa), sb_u = tc::smem_u32(sb);
  uint32_t phase[2] = {0, 0};
  bool failed = false;

  auto issue_tile = [&](int t) {  // thread 0 only
    tc::fence_after_sync();
    const uint32_t acc = tmem + (uint32_t)(t & 1) * N;
    bool first = true;
#pragma unroll
    for (int tap = 0; tap < 9; tap++) {
      const uint32_t off = (uint32_t)((tap / 3) * IDIM + tap % 3 + t * 128) * 16;
#pragma unroll
      for (int s = 0; s < C / 32; s++) {
        const uint64_t ad = tc::make_desc(sa_u + (uint32_t)(2 * s) * Cfg::PLANE + off, Cfg::PLANE, 128);
        const uint64_t bd = tc::make_desc(sb_u + (uint32_t)((tap * Cfg::KC + 2 * s) * N) * 16, N * 16, 128);
        tc::mma_i8(acc, ad, bd, idesc, first ? 0u : 1u);
        first = false;
      }
    }
    tc::commit(&bar[t & 1]);
  };

  const int ngroups = (n_img + G - 1) / G;
  for (int grp = blockIdx.x; grp < ngroups; grp += gridDim.x) {
    const int img0 = grp * G;
    // ---- 1. expand bit-packed activations to int8 operand planes ---------------------------------------------------------
    constexpr int CW = C / 32;
    for (int i = tid; i < Cfg::ROWS * CW; i += 128) {
      const int j = i % CW, m = i / CW;            // 32-channel word j of concatenated pixel m
      const int gi = m / Cfg::PXI;
      uint32_t w0 = 0, w1 = 0;
      if (img0 + gi < n_img) {
        const size_t base = ((size_t)(img0 + gi) * Cfg::PXI + (m - gi * Cfg::PXI)) * (CW * Cfg::IPL);
        w0 = __ldg(in + base + j);
        if (INM == IN_TERN) w1 = __ldg(in + base + CW + j);
      }
#pragma unroll
      for (int h = 0; h < 2; h++) {                // two 16-channel chunks per word
        uint32_t v[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const uint32_t a = spread4((w0 >> (16 * h + 4 * q)) & 0xf);
          if (INM == IN_TERN) {
            const uint32_t p = spread4((w1 >> (16 * h + 4 * q)) & 0xf);
            v[q] = (a * 0xffu) & ~(p * 0xfeu);    // nz ? (pos ? +1 : -1) : 0
          } else {
            v[q] = a;                              // {0,1}
          }
        }
        *reinterpret_cast<uint4*>(sa + (size_t)(2 * j + h) * Cfg::PLANE + (size_t)m * 16) = make_uint4(v[0], v[1], v[2], v[3]);
      }
    }
    tc::fence_async_smem();
    __syncthreads();
    // ---- 2./3. tiles: MMA issue one tile ahead of the epilogue -------------------------------------------------------------
    if (tid == 0) issue_tile(0);
    for (int t = 0; t < Cfg::TILES; t++) {
      if (tid == 0 && t + 1 < Cfg::TILES) issue_tile(t + 1);
      if (!failed && !tc::mbar_wait(&bar[t & 1], phase[t & 1])) failed = true;
      phase[t & 1] ^= 1;
      tc::fence_after_sync();
      const int m = t * 128 + warp * 32 + lane;
      const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(t & 1) * N;
#pragma unroll
      for (int j = 0; j < Cfg::NW; j++) {
        uint32_t r[32];
        tc::tmem_ld32(trow + j * 32, r);
        tc::tmem_ld_wait();
        if (tap) {  // parity taps: the int32 MMA sums of every valid output pixel
          const int gi = m / Cfg::PXI, rem = m - gi * Cfg::PXI, y = rem / IDIM, x = rem - y * IDIM;
          if (m < Cfg::ROWS && y < OD && x < OD && img0 + gi < n_img) {
#pragma unroll
            for (int c = 0; c < 32; c++) tap[((size_t)(img0 + gi) * (OD * OD) + y * OD + x) * N + j * 32 + c] = (int)r[c];
          }
        }
        uint32_t b0 = 0, b1 = 0;
#pragma unroll
        for (int c = 0; c < 32; c++) {
          const int S = (int)r[c];
          const int t0 = sthr[j * 32 + c] < S;
          if (OUTM == OUT_BIT) {
            b0 |= (uint32_t)t0 << c;
          } else {
            const int t1 = sthr[N + j * 32 + c] < S;
            b0 |= (uint32_t)(1 ^ t0 ^ t1) << c;   // nz
            b1 |= (uint32_t)(t0 & t1) << c;       // pos
          }
        }
        sbits[(size_t)m * (Cfg::NW * Cfg::OPL) + j] = b0;
        if (OUTM == OUT_TERN) sbits[(size_t)m * (Cfg::NW * Cfg::OPL) + Cfg::NW + j] = b1;
      }
      tc::fence_before_sync();
      __syncthreads();
    }
    // ---- 4. pool + store ----------------------------------------------------------------------------------------------------
    constexpr int WPP = Cfg::NW * Cfg::OPL;  // words per output pixel
    for (int i = tid; i < G * Cfg::PD * Cfg::PD * Cfg::NW; i += 128) {
      const int j = i % Cfg::NW, op = (i / Cfg::NW) % (Cfg::PD * Cfg::PD), gi = i / (Cfg::NW * Cfg::PD * Cfg::PD);
      if (img0 + gi >= n_img) continue;
      const int oy = op / Cfg::PD, ox = op % Cfg::PD;
      uint32_t o0, o1 = 0;
      if (POOL) {
        const uint32_t* r00 = sbits + (size_t)(gi * Cfg::PXI + (2 * oy) * IDIM + 2 * ox) * WPP;
        const uint32_t* r01 = r00 + WPP;
        const uint32_t* r10 = r00 + (size_t)IDIM * WPP;
        const uint32_t* r11 = r10 + WPP;
        if (OUTM == OUT_BIT) {
          o0 = r00[j] | r01[j] | r10[j] | r11[j];                       // maxpool.h:77-83
        } else {                                                         // signed max, init -1 (maxpool.h:141,156-160)
          const uint32_t P = r00[Cfg::NW + j] | r01[Cfg::NW + j] | r10[Cfg::NW + j] | r11[Cfg::NW + j];
          const uint32_t M = (r00[j] & ~r00[Cfg::NW + j]) & (r01[j] & ~r01[Cfg::NW + j]) & (r10[j] & ~r10[Cfg::NW + j]) & (r11[j] & ~r11[Cfg::NW + j]);
          o0 = P | M; o1 = P;
        }
      } else {
        const uint32_t* r = sbits + (size_t)(gi * Cfg::PXI + oy * IDIM + ox) * WPP;
        o0 = r[j];
        if (OUTM == OUT_TERN) o1 = r[Cfg::NW + j];
      }
      uint32_t* dst = out + ((size_t)(img0 + gi) * (Cfg::PD * Cfg::PD) + op) * WPP;
      dst[j] = o0;
      if (OUTM == OUT_TERN) dst[Cfg::NW + j] = o1;
    }
    __syncthreads();  // sbits / sa are rewritten by the next group
  }
  if (failed && status) *status = 1;
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, Cfg::TMEM_COLS);
}

template <int C, int N, int IDIM, int OD, bool POOL, int G, int INM, int OUTM>
static bool launch_cfg(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int* status, int32_t* tap, cudaStream_t s) {
  using Cfg = ConvTcCfg<C, N, IDIM, OD, POOL, G, INM, OUTM>;
  static_assert(Cfg::SMEM <= 226 * 1024, "shared memory budget");
  auto kern = k_conv_tc<C, N, IDIM, OD, POOL, G, INM, OUTM>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
  int per_sm = (227 * 1024) / (Cfg::SMEM + 1024);
  if (per_sm * Cfg::TMEM_COLS > 512) per_sm = 512 / Cfg::TMEM_COLS;
  if (per_sm < 1) per_sm = 1;
  const int ngroups = (n_img + G - 1) / G;
  int grid = num_sms * per_sm;
  if (grid > ngroups) grid = ngroups;
  kern<<<grid, 128, Cfg::SMEM, s>>>(L, in, out, n_img, status, tap);
  return true;
}

// Returns false if this layer shape has no tensor-core instantiation (the caller then uses the LOP3+POPC kernel).
bool launch_conv_tc(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int* status, int32_t* tap, cudaStream_t s) {
  const LayerDesc& d = L.d;
  if (d.kind != KIND_CONV || !L.w8 || !L.thr_tc || n_img <= 0) return false;
  const bool a1 = d.in_mode == IN_XNOR && d.out_mode == OUT_BIT, a2 = d.in_mode == IN_TERN && d.out_mode == OUT_TERN;
  if (d.ifm_ch == 64 && d.N == 64 && d.ifm_dim == 30 && d.pool) {
    if (a1) return launch_cfg<64, 64, 30, 28, true, 1, IN_XNOR, OUT_BIT>(L, in, out, n_img, num_sms, status, tap, s);
    if (a2) return launch_cfg<64, 64, 30, 28, true, 1, IN_TERN, OUT_TERN>(L, in, out, n_img, num_sms, status, tap, s);
  }
  if (d.ifm_ch == 64 && d.N == 128 && d.ifm_dim == 14 && !d.pool) {
    if (a1) return launch_cfg<64, 128, 14, 12, false, 6, IN_XNOR, OUT_BIT>(L, in, out, n_img, num_sms, status, tap, s);
    if (a2) return launch_cfg<64, 128, 14, 12, false, 4, IN_TERN, OUT_TERN>(L, in, out, n_img, num_sms, status, tap, s);
  }
  if (d.ifm_ch == 128 && d.N == 128 && d.ifm_dim == 12 && d.pool) {
    if (a1) return launch_cfg<128, 128, 12, 10, true, 2, IN_XNOR, OUT_BIT>(L, in, out, n_img, num_sms, status, tap, s);
    if (a2) return launch_cfg<128, 128, 12, 10, true, 2, IN_TERN, OUT_TERN>(L, in, out, n_img, num_sms, status, tap, s);
  }
  return false;
}

}  // namespace bnnb
