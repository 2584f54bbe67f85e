// Pipe micro-benchmarks: the measured denominators for "achieved binary-op/s against the integer-pipe or tensor-pipe peak".
#include <cuda_runtime.h>
#include <cstdint>
#include "../../include/bnn_b200.h"

namespace {
constexpr int kIters = 4096, kUnroll = 8;

__global__ void __launch_bounds__(512) mb_popc(uint32_t* out, long long* cyc) {
  uint32_t a[kUnroll], acc[kUnroll];
#pragma unroll
  for (int i = 0; i < kUnroll; i++) { a[i] = threadIdx.x * 2654435761u + i; acc[i] = 0; }
  const long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int i = 0; i < kUnroll; i++) { acc[i] += __popc(a[i]); a[i] += acc[i]; }  // dependent chains, kUnroll independent
  }
  const long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kUnroll; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void __launch_bounds__(512) mb_lop3(uint32_t* out, long long* cyc) {
  uint32_t a[kUnroll], b = threadIdx.x | 1, c = blockIdx.x * 7 + 3;
#pragma unroll
  for (int i = 0; i < kUnroll; i++) a[i] = threadIdx.x * 2654435761u + i;
  const long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int i = 0; i < kUnroll; i++) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
  }
  const long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kUnroll; i++) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
__global__ void __launch_bounds__(512) mb_imma(uint32_t* out, long long* cyc) {
  int d[kUnroll][4];
  uint32_t a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = threadIdx.x * 5, a3 = threadIdx.x * 7, b0 = blockIdx.x, b1 = blockIdx.x * 11;
#pragma unroll
  for (int i = 0; i < kUnroll; i++) d[i][0] = d[i][1] = d[i][2] = d[i][3] = 0;
  const long long t0 = clock64();
  for (int it = 0; it < kIters; it++) {
#pragma unroll
    for (int i = 0; i < kUnroll; i++)
      asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+r"(d[i][0]), "+r"(d[i][1]), "+r"(d[i][2]), "+r"(d[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  const long long t1 = clock64();
  int s = 0;
#pragma unroll
  for (int i = 0; i < kUnroll; i++) s += d[i][0] + d[i][1] + d[i][2] + d[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = (uint32_t)s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
}  // namespace

extern "C" double bnnb_microbench(int which, double* clock_mhz) {
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) return -1;
  cudaSetDevice(0);
  const int blocks = prop.multiProcessorCount * 4, threads = 512;
  uint32_t* out; long long* cyc;
  if (cudaMalloc(&out, (size_t)blocks * threads * 4) != cudaSuccess) return -1;
  cudaMalloc(&cyc, blocks * 8);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    if (which == 0) mb_popc<<<blocks, threads>>>(out, cyc);
    else if (which == 1) mb_imma<<<blocks, threads>>>(out, cyc);
    else mb_lop3<<<blocks, threads>>>(out, cyc);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  long long h = 0;
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const bool ok = cudaGetLastError() == cudaSuccess;
  cudaFree(out); cudaFree(cyc); cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (!ok) return -1;
  // a block's cycle count over the kernel's wall time bounds the clock from below (blocks run 2 waves at most)
  if (clock_mhz) *clock_mhz = (double)h / (best * 1e-3) / 1e6;
  const double per_thread = (double)kIters * kUnroll;
  const double total = per_thread * blocks * threads;
  if (which == 1) return total / 32.0 * (16.0 * 8 * 32) / (best * 1e-3);  // one mma per warp = 4096 MACs
  return total / (best * 1e-3);
}
