// tcgen05 kind::i8 probe: (1) correctness of the descriptor encodings used by the conv kernels (K-major, no swizzle, operand
// start shifted by whole 16-byte rows = the implicit-im2col trick), (2) issue-rate measurement for M=128, N in {64,128,256}.
#include <cuda_runtime.h>
#include <cstdint>
#include "../../include/bnn_b200.h"
#include "tcgen05.cuh"

namespace {
using namespace bnnb::tc;

// A: [rows_a][K] int8 row-major (rows_a >= 128 + shift), B: [N][K] int8 row-major, D: [128][N] int32.
// D[m][n] = sum_k A[m + shift][k] * B[n][k].  reps > 1 re-issues the same MMAs (timing); cycles returned in *cyc.
__global__ void __launch_bounds__(128) k_tc_probe(const int8_t* __restrict__ A, const int8_t* __restrict__ B, int32_t* __restrict__ D,
                                                   int rows_a, int N, int K, int shift, int reps, long long* cyc, int* status) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int kch = K / 16;                       // 16-byte K chunks
  const uint32_t plane_a = (uint32_t)rows_a * 16, plane_b = (uint32_t)N * 16;
  uint8_t* sa = smem;
  uint8_t* sb = smem + (size_t)kch * plane_a;
  for (int i = threadIdx.x; i < rows_a * kch; i += blockDim.x) {
    const int r = i / kch, g = i % kch;
    *reinterpret_cast<uint4*>(sa + (size_t)g * plane_a + r * 16) = *reinterpret_cast<const uint4*>(A + (size_t)r * K + g * 16);
  }
  for (int i = threadIdx.x; i < N * kch; i += blockDim.x) {
    const int r = i / kch, g = i % kch;
    *reinterpret_cast<uint4*>(sb + (size_t)g * plane_b + r * 16) = *reinterpret_cast<const uint4*>(B + (size_t)r * K + g * 16);
  }
  fence_async_smem();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc(&tmem_base, 256);
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init_fence(); }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_base;
  const uint32_t idesc = make_idesc_i8(128, N);
  long long t0 = 0, t1 = 0;
  if (threadIdx.x == 0) {
    t0 = clock64();
    for (int rep = 0; rep < reps; rep++)
      for (int ks = 0; ks < K / 32; ks++) {
        const uint64_t ad = make_desc(smem_u32(sa) + (uint32_t)(2 * ks) * plane_a + (uint32_t)shift * 16, plane_a, 128);
        const uint64_t bd = make_desc(smem_u32(sb) + (uint32_t)(2 * ks) * plane_b, plane_b, 128);
        mma_i8(tmem, ad, bd, idesc, (ks > 0 || rep > 0) ? 1u : 0u);
      }
    commit(&bar);
  }
  const bool ok = mbar_wait(&bar, 0);
  if (threadIdx.x == 0) { t1 = clock64(); *cyc = t1 - t0; }
  fence_after_sync();
  if (!ok) { if (threadIdx.x == 0) *status = 1; }
  else {
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; j++) D[(size_t)(warp * 32 + lane) * N + c0 + j] = (int32_t)r[j];
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}
}  // namespace

// Host buffers in, host buffer out.  Returns 0, 1 (mbarrier timeout) or -1 (CUDA error); *cycles = issue-to-completion SM cycles.
extern "C" int bnnb_tc_probe(const int8_t* A, const int8_t* B, int32_t* D, int rows_a, int N, int K, int shift, int reps, long long* cycles) {
  if (N % 16 || N < 16 || N > 256 || K % 32 || rows_a < 128 + shift || rows_a % 8) return -1;
  const size_t smem = (size_t)(K / 16) * (rows_a + N) * 16;
  if (smem > 200 * 1024) return -1;
  int8_t *dA, *dB; int32_t* dD; long long* dC; int* dS;
  if (cudaMalloc(&dA, (size_t)rows_a * K) != cudaSuccess) return -1;
  cudaMalloc(&dB, (size_t)N * K); cudaMalloc(&dD, (size_t)128 * N * 4); cudaMalloc(&dC, 8); cudaMalloc(&dS, 4);
  cudaMemcpy(dA, A, (size_t)rows_a * K, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B, (size_t)N * K, cudaMemcpyHostToDevice);
  cudaMemset(dS, 0, 4);
  cudaMemset(dD, 0xff, (size_t)128 * N * 4);
  cudaFuncSetAttribute(k_tc_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  k_tc_probe<<<1, 128, smem>>>(dA, dB, dD, rows_a, N, K, shift, reps, dC, dS);
  cudaError_t e = cudaDeviceSynchronize();
  int st = 0;
  if (e == cudaSuccess) {
    cudaMemcpy(D, dD, (size_t)128 * N * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(cycles, dC, 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(&st, dS, 4, cudaMemcpyDeviceToHost);
  }
  cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dC); cudaFree(dS);
  return e == cudaSuccess ? st : -1;
}

// ---- fp4 probe: A [rows_a][K/2] packed e2m1, B [N][K/2], D[m][n] = sum_k A[m+shift][k] * B[n][k] (fp32 -> int32) --------------
namespace {
__global__ void __launch_bounds__(128) k_tc_probe_fp4(const uint8_t* __restrict__ A, const uint8_t* __restrict__ B, int32_t* __restrict__ D,
                                                       int rows_a, int N, int K, int shift, int reps, long long* cyc, int* status, const float* __restrict__ bias, int mix) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int kch = K / 32;  // 16-byte chunks (32 fp4 each)
  const uint32_t plane_a = (uint32_t)rows_a * 16, plane_b = (uint32_t)N * 16;
  uint8_t* sa = smem;
  uint8_t* sb = smem + (size_t)kch * plane_a;
  for (int i = threadIdx.x; i < rows_a * kch; i += blockDim.x) {
    const int r = i / kch, g = i % kch;
    *reinterpret_cast<uint4*>(sa + (size_t)g * plane_a + r * 16) = *reinterpret_cast<const uint4*>(A + (size_t)r * (K / 2) + g * 16);
  }
  for (int i = threadIdx.x; i < N * kch; i += blockDim.x) {
    const int r = i / kch, g = i % kch;
    *reinterpret_cast<uint4*>(sb + (size_t)g * plane_b + r * 16) = *reinterpret_cast<const uint4*>(B + (size_t)r * (K / 2) + g * 16);
  }
  fence_async_smem();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc(&tmem_base, 512);
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init_fence(); }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_base;
  const uint32_t sf = tmem + 256;  // 16 columns x 128 lanes of ue8m0 1.0 serve as every scale factor
  tmem_fill16(sf + ((uint32_t)(warp * 32) << 16), 0x7f7f7f7fu);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  if (bias) {  // accumulator pre-loaded with bias[n] in every row; all MMAs then accumulate; D returns raw fp32 bits
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t r[32];
#pragma unroll
      for (int j = 0; j < 32; j++) r[j] = __float_as_uint(bias[c0 + j]);
      tmem_st32(tmem + ((uint32_t)(warp * 32) << 16) + c0, r);
    }
    tmem_st_wait();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
  }
  const uint32_t idesc = make_idesc_mxf4(128, N);
  long long t0 = 0, t1 = 0;
  if (threadIdx.x == 0 && mix != 0) {
    // timing experiment (results are not meaningful): `reps` groups of |mix| fp4 MMAs, each group followed by one kind::i8 MMA
    // (mix > 0) or by one more fp4 MMA (mix < 0, the control), all on constant descriptors so that the issue loop is lean
    const uint64_t ad = make_desc(smem_u32(sa), plane_a, 128), bd = make_desc(smem_u32(sb), plane_b, 128);
    const uint32_t id8 = make_idesc_i8(128, 64);
    const int grp = mix > 0 ? mix : -mix;
    t0 = clock64();
    for (int rep = 0; rep < reps; rep++) {
      for (int i = 0; i < grp; i++) mma_mxf4(tmem, ad, bd, idesc, sf, sf, 1u);
      if (mix > 0) mma_i8(tmem + 288, ad, bd, id8, 1u);
      else mma_mxf4(tmem + 288, ad, bd, idesc, sf, sf, 1u);
    }
    commit(&bar);
  } else if (threadIdx.x == 0) {
    t0 = clock64();
    for (int rep = 0; rep < reps; rep++)
      for (int ks = 0; ks < K / 64; ks++) {
        const uint64_t ad = make_desc(smem_u32(sa) + (uint32_t)(2 * ks) * plane_a + (uint32_t)shift * 16, plane_a, 128);
        const uint64_t bd = make_desc(smem_u32(sb) + (uint32_t)(2 * ks) * plane_b, plane_b, 128);
        mma_mxf4(tmem, ad, bd, idesc, sf, sf, (ks > 0 || rep > 0 || bias) ? 1u : 0u);
      }
    commit(&bar);
  }
  const bool ok = mbar_wait(&bar, 0);
  if (threadIdx.x == 0) { t1 = clock64(); *cyc = t1 - t0; }
  fence_after_sync();
  if (!ok) { if (threadIdx.x == 0) *status = 1; }
  else {
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t r[32];
      tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, r);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; j++) D[(size_t)(warp * 32 + lane) * N + c0 + j] = bias ? (int32_t)r[j] : __float2int_rn(__uint_as_float(r[j]));
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}
}  // namespace

extern "C" int bnnb_tc_probe_fp4_mix(const uint8_t* A, const uint8_t* B, int32_t* D, int rows_a, int N, int K, int shift, int reps, long long* cycles,
                                     const float* bias, int mix);
extern "C" int bnnb_tc_probe_fp4_bias(const uint8_t* A, const uint8_t* B, int32_t* D, int rows_a, int N, int K, int shift, int reps, long long* cycles,
                                      const float* bias) {
  return bnnb_tc_probe_fp4_mix(A, B, D, rows_a, N, K, shift, reps, cycles, bias, 0);
}
extern "C" int bnnb_tc_probe_fp4(const uint8_t* A, const uint8_t* B, int32_t* D, int rows_a, int N, int K, int shift, int reps, long long* cycles) {
  return bnnb_tc_probe_fp4_bias(A, B, D, rows_a, N, K, shift, reps, cycles, nullptr);
}
extern "C" int bnnb_tc_probe_fp4_mix(const uint8_t* A, const uint8_t* B, int32_t* D, int rows_a, int N, int K, int shift, int reps, long long* cycles,
                                     const float* bias, int mix) {
  if (N % 32 || N < 32 || N > 256 || K % 64 || rows_a < 128 + shift || rows_a % 8) return -1;
  const size_t smem = (size_t)(K / 32) * (rows_a + N) * 16;
  if (smem > 200 * 1024) return -1;
  uint8_t *dA, *dB; int32_t* dD; long long* dC; int* dS;
  if (cudaMalloc(&dA, (size_t)rows_a * K / 2) != cudaSuccess) return -1;
  cudaMalloc(&dB, (size_t)N * K / 2); cudaMalloc(&dD, (size_t)128 * N * 4); cudaMalloc(&dC, 8); cudaMalloc(&dS, 4);
  cudaMemcpy(dA, A, (size_t)rows_a * K / 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, B, (size_t)N * K / 2, cudaMemcpyHostToDevice);
  float* dBias = nullptr;
  if (bias) { cudaMalloc(&dBias, (size_t)N * 4); cudaMemcpy(dBias, bias, (size_t)N * 4, cudaMemcpyHostToDevice); }
  cudaMemset(dS, 0, 4);
  cudaMemset(dD, 0xff, (size_t)128 * N * 4);
  cudaFuncSetAttribute(k_tc_probe_fp4, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  k_tc_probe_fp4<<<1, 128, smem>>>(dA, dB, dD, rows_a, N, K, shift, reps, dC, dS, dBias, mix);
  cudaError_t e = cudaDeviceSynchronize();
  int st = 0;
  if (e == cudaSuccess) {
    cudaMemcpy(D, dD, (size_t)128 * N * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(cycles, dC, 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(&st, dS, 4, cudaMemcpyDeviceToHost);
  }
  if (dBias) cudaFree(dBias);
  cudaFree(dA); cudaFree(dB); cudaFree(dD); cudaFree(dC); cudaFree(dS);
  return e == cudaSuccess ? st : -1;
}

// ---- lean issue-rate probe: fully unrolled kind::mxf4 MMAs on (lo, hi) descriptor words, converged warp + elect_one, so that the
// measured clocks per MMA are the tensor pipe's and not the issuing thread's.  taps != 0: the A start moves through the nine
// (ky*30 + kx) row offsets of a 3x3 window (16-byte, mostly not 128-byte aligned), as in the implicit-im2col kernels. -----------------
namespace {
__global__ void __launch_bounds__(128) k_tc_rate_fp4(int N, int taps, int iters, long long* cyc, int* status) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (200 * 1024) / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0x22222222u, 0x0a0a0a0au, 0x20202020u, 0x02020202u);
  fence_async_smem();
  if (warp == 0) tmem_alloc(&tmem_base, 512);
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init_fence(); }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem = tmem_base, sf = tmem + 384;
  tmem_fill16(sf + ((uint32_t)(warp * 32) << 16), 0x7f7f7f7fu);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  long long t0 = 0;
  if (warp == 0) {
    const bool leader = elect_one();
    const uint32_t idesc = make_idesc_mxf4(128, N), d_hi = desc_hi(128);
    const uint32_t a_lo = desc_lo(smem_u32(smem), 1024 * 16), b_lo = desc_lo(smem_u32(smem) + 64 * 1024, (uint32_t)N * 16);
    t0 = clock64();
    for (int it = 0; it < iters; it++) {
      if (leader) {
#pragma unroll
        for (int tap = 0; tap < 9; tap++) {
          const uint32_t off = taps ? (uint32_t)((tap / 3) * 30 + tap % 3) : 0u;
          mma_mxf4_lh(tmem + (uint32_t)(it & 1) * 128, a_lo + off, d_hi, b_lo + (uint32_t)(tap * 2 * N), d_hi, idesc, sf, sf, 1u);
        }
      }
      __syncwarp();
    }
    if (leader) commit(&bar);
    __syncwarp();
  }
  const bool ok = mbar_wait(&bar, 0);
  if (threadIdx.x == 0) { *cyc = clock64() - t0; if (!ok) *status = 1; }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}
}  // namespace

// clocks per MMA (M=128, K=64, fp4) at width N with 9*iters MMAs issued back to back
extern "C" double bnnb_tc_rate_fp4(int N, int taps) {
  if (N % 32 || N < 32 || N > 128) return -1;
  long long* dC; int* dS;
  if (cudaMalloc(&dC, 8) != cudaSuccess) return -1;
  cudaMalloc(&dS, 4);
  cudaMemset(dS, 0, 4);
  cudaFuncSetAttribute(k_tc_rate_fp4, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 512;
  long long c[2] = {0, 0};
  for (int r = 0; r < 2; r++) {
    k_tc_rate_fp4<<<1, 128, 200 * 1024>>>(N, taps, r == 0 ? 1 : iters, dC, dS);
    if (cudaDeviceSynchronize() != cudaSuccess) { cudaFree(dC); cudaFree(dS); return -1; }
    cudaMemcpy(&c[r], dC, 8, cudaMemcpyDeviceToHost);
  }
  int st = 0;
  cudaMemcpy(&st, dS, 4, cudaMemcpyDeviceToHost);
  cudaFree(dC); cudaFree(dS);
  return st ? -1 : (double)(c[1] - c[0]) / (9.0 * (iters - 1));
}
