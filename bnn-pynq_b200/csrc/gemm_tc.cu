// K-streamed tensor-core MVTU for the layers whose weights do not fit in shared memory next to their activations
// (CNV layers 4..8, LFC layers): rows = (image, output pixel), gathered im2col, tcgen05.mma kind::i8, TMEM accumulators.
//
// One CTA = one 128-row tile x one slice of <= 256 output channels.  K is walked in 64-byte slabs (two K=32 MMAs):
// the slab of the weight matrix (N x 64 int8, from the re-packed [N][K] array, L2-resident) and the slab of the activation
// rows (expanded from the bit-packed tensor to int8) are written to one of two shared-memory buffers in the UMMA K-major
// un-swizzled layout; the MMAs of slab s run while slab s+1 is being staged (an mbarrier per buffer, committed by the MMAs,
// says when it may be overwritten).  Epilogue as in conv_tc.cu: one row per thread, integer threshold, bits packed in
// registers, direct global store (or 16-bit scores for the last CNV layer).
// Replaces StreamingFCLayer_Batch / ConvLayer_Batch + ThresholdsActivation / PassThroughActivation for these layers
// (fclayer.h:83-111, convlayer.h:85-121, activations.hpp:85-148).
#include <climits>
#include "kernels.cuh"
#include "tcgen05.cuh"

namespace bnnb {
namespace {

__device__ __forceinline__ uint32_t spread4g(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

template <int N, int INM, int OUTM>  // N: channels of this CTA's slice (64, 128 or 256)
__global__ void __launch_bounds__(128) k_gemm_tc(DevLayer L, const uint32_t* __restrict__ in, uint32_t* __restrict__ out, long long rows, int32_t* __restrict__ tap,
                                                  int* __restrict__ status) {
  constexpr int A_SLAB = 4 * 128 * 16, B_SLAB = 4 * N * 16, BUF = A_SLAB + B_SLAB;
  constexpr int IPL = INM == IN_TERN ? 2 : 1;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar[2];
  __shared__ uint32_t tmem_base;
  __shared__ int sthr[2 * N];
  const LayerDesc& d = L.d;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = blockIdx.y * N;
  const long long row0 = (long long)blockIdx.x * 128;
  for (int i = tid; i < 2 * N; i += 128) sthr[(i & 1) * N + (i >> 1)] = L.thr_tc[2 * n0 + i];  // [2][N]
  if (warp == 0) tc::tmem_alloc(&tmem_base, N < 32 ? 32 : N);
  if (tid == 0) { tc::mbar_init(&bar[0], 1); tc::mbar_init(&bar[1], 1); tc::mbar_init_fence(); }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_base, idesc = tc::make_idesc_i8(128, N);
  uint32_t ph[2] = {0, 0};
  bool failed = false;

  // this thread's row: source pixel base for the im2col gather
  const int opx = d.ofm_dim * d.ofm_dim, C = d.ifm_ch, CW = C / 32;
  const long long r = row0 + tid;
  const bool rvalid = r < rows;
  const long long img = rvalid ? r / opx : 0;
  const int op = rvalid ? (int)(r % opx) : 0, oy = op / d.ofm_dim, ox = op % d.ofm_dim;
  const size_t in_img_words = (size_t)d.ifm_dim * d.ifm_dim * CW * IPL;
  const int nslab = d.K / 64;

  for (int s = 0; s < nslab; s++) {
    uint8_t* sa = smem + (size_t)(s & 1) * BUF;
    uint8_t* sb = sa + A_SLAB;
    if (s >= 2) {  // the MMAs that read this buffer two slabs ago must have completed
      if (!failed && !tc::mbar_wait(&bar[s & 1], ph[s & 1])) failed = true;
      ph[s & 1] ^= 1;
    }
    // weights: rows n0..n0+N, K bytes [64s, 64s+64) -> 4 chunk planes of N x 16 bytes
    for (int i = tid; i < N * 4; i += 128) {
      const int n = i >> 2, g = i & 3;
      *reinterpret_cast<uint4*>(sb + ((size_t)g * N + n) * 16) =
          __ldg(reinterpret_cast<const uint4*>(L.w8 + (size_t)(n0 + n) * L.kpad + (size_t)s * 64 + g * 16));
    }
    // activations: this thread's row, 64 channels of one tap
    {
      const int k0 = s * 64, tap = k0 / C, c0 = k0 % C;
      uint32_t w0[2] = {0, 0}, w1[2] = {0, 0};
      if (rvalid) {
        const int sp = (oy + tap / 3) * d.ifm_dim + ox + tap % 3;
        const uint32_t* src = in + img * in_img_words + (size_t)sp * CW * IPL + c0 / 32;
        w0[0] = __ldg(src); w0[1] = __ldg(src + 1);
        if (INM == IN_TERN) { w1[0] = __ldg(src + CW); w1[1] = __ldg(src + CW + 1); }
      }
#pragma unroll
      for (int g = 0; g < 4; g++) {
        uint32_t v[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const uint32_t a = spread4g((w0[g >> 1] >> (16 * (g & 1) + 4 * q)) & 0xf);
          if (INM == IN_TERN) { const uint32_t p = spread4g((w1[g >> 1] >> (16 * (g & 1) + 4 * q)) & 0xf); v[q] = (a * 0xffu) & ~(p * 0xfeu); }
          else v[q] = a;
        }
        *reinterpret_cast<uint4*>(sa + ((size_t)g * 128 + tid) * 16) = make_uint4(v[0], v[1], v[2], v[3]);
      }
    }
    tc::fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      tc::fence_after_sync();
      const uint32_t sa_u = tc::smem_u32(sa), sb_u = tc::smem_u32(sb);
      tc::mma_i8(tmem, tc::make_desc(sa_u, 128 * 16, 128), tc::make_desc(sb_u, N * 16, 128), idesc, s > 0 ? 1u : 0u);
      tc::mma_i8(tmem, tc::make_desc(sa_u + 2 * 128 * 16, 128 * 16, 128), tc::make_desc(sb_u + 2 * N * 16, N * 16, 128), idesc, 1u);
      tc::commit(&bar[s & 1]);
    }
  }
  // drain: the last one or two commits
  for (int s = (nslab >= 2 ? nslab - 2 : 0); s < nslab; s++) {
    if (!failed && !tc::mbar_wait(&bar[s & 1], ph[s & 1])) failed = true;
    ph[s & 1] ^= 1;
  }
  tc::fence_after_sync();
  // ---- epilogue: row = row0 + warp*32 + lane -------------------------------------------------------------------------------
  const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
  constexpr int NW = N / 32;
  const int nwt = d.N / 32;  // words per plane of the whole layer
#pragma unroll 1
  for (int j = 0; j < NW; j++) {
    uint32_t v[32];
    tc::tmem_ld32(trow + j * 32, v);
    tc::tmem_ld_wait();
    if (tap && rvalid) {  // parity taps: the int32 MMA sums of this row
#pragma unroll
      for (int c = 0; c < 32; c++) tap[(size_t)r * d.N + n0 + j * 32 + c] = (int)v[c];
    }
    if (OUTM == OUT_SCORE) {  // PassThroughActivation<ap_uint<16>>: (S + offset) mod 2^16, offset = #negative weights for 1-bit inputs
      if (rvalid) {
        int16_t* o16 = reinterpret_cast<int16_t*>(out) + (size_t)r * d.N + n0 + j * 32;
#pragma unroll
        for (int c = 0; c < 32; c += 2) {
          const uint32_t lo = (uint32_t)((int)v[c] + sthr[j * 32 + c]) & 0xffff, hi = (uint32_t)((int)v[c + 1] + sthr[j * 32 + c + 1]) & 0xffff;
          *reinterpret_cast<uint32_t*>(o16 + c) = lo | (hi << 16);
        }
      }
    } else {
      uint32_t b0 = 0, b1 = 0;
#pragma unroll
      for (int c = 0; c < 32; c++) {
        const int S = (int)v[c];
        const int t0 = sthr[j * 32 + c] < S;
        if (OUTM == OUT_BIT) b0 |= (uint32_t)t0 << c;
        else { const int t1 = sthr[N + j * 32 + c] < S; b0 |= (uint32_t)(1 ^ t0 ^ t1) << c; b1 |= (uint32_t)(t0 & t1) << c; }
      }
      if (rvalid) {
        uint32_t* dst = out + (size_t)r * nwt * (OUTM == OUT_TERN ? 2 : 1) + n0 / 32 + j;
        dst[0] = b0;
        if (OUTM == OUT_TERN) dst[nwt] = b1;
      }
    }
  }
  if (failed && status) *status = 1;
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, N < 32 ? 32 : N);
}

template <int N, int INM, int OUTM>
void launch_t(const DevLayer& L, const uint32_t* in, uint32_t* out, long long rows, int* status, int32_t* tap, cudaStream_t s) {
  constexpr int SMEM = 2 * (4 * 128 * 16 + 4 * N * 16);
  cudaFuncSetAttribute(k_gemm_tc<N, INM, OUTM>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
  dim3 grid((unsigned)((rows + 127) / 128), (unsigned)(L.d.N / N));
  k_gemm_tc<N, INM, OUTM><<<grid, 128, SMEM, s>>>(L, in, out, rows, tap, status);
}

template <int N>
bool launch_n(const DevLayer& L, const uint32_t* in, uint32_t* out, long long rows, int* status, int32_t* tap, cudaStream_t s) {
  const int im = L.d.in_mode, om = L.d.out_mode;
  if (im == IN_XNOR && om == OUT_BIT) launch_t<N, IN_XNOR, OUT_BIT>(L, in, out, rows, status, tap, s);
  else if (im == IN_XNOR && om == OUT_SCORE) launch_t<N, IN_XNOR, OUT_SCORE>(L, in, out, rows, status, tap, s);
  else if (im == IN_TERN && om == OUT_TERN) launch_t<N, IN_TERN, OUT_TERN>(L, in, out, rows, status, tap, s);
  else if (im == IN_TERN && om == OUT_SCORE) launch_t<N, IN_TERN, OUT_SCORE>(L, in, out, rows, status, tap, s);
  else if (im == IN_TERN && om == OUT_BIT) launch_t<N, IN_TERN, OUT_BIT>(L, in, out, rows, status, tap, s);
  else return false;
  return true;
}

}  // namespace

// Returns false when the layer cannot run on this kernel (the caller falls back to the LOP3+POPC kernel).
bool launch_gemm_tc(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int* status, int32_t* tap, cudaStream_t s) {
  const LayerDesc& d = L.d;
  if (!L.w8 || !L.thr_tc || n_img <= 0 || d.pool || d.K % 64 || d.ifm_ch % 64) return false;
  if (d.in_mode != IN_XNOR && d.in_mode != IN_TERN) return false;
  const long long rows = (long long)n_img * d.ofm_dim * d.ofm_dim;
  const long long tiles = (rows + 127) / 128;
  // widest channel slice that still gives every SM about two CTAs (wide slices re-use the expanded activation slab more)
  if (d.N % 256 == 0 && tiles * (d.N / 256) >= 2LL * num_sms) return launch_n<256>(L, in, out, rows, status, tap, s);
  if (d.N % 128 == 0 && tiles * (d.N / 128) >= 2LL * num_sms) return launch_n<128>(L, in, out, rows, status, tap, s);
  if (d.N % 64 == 0) return launch_n<64>(L, in, out, rows, status, tap, s);
  return false;
}

}  // namespace bnnb
