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
// generic ALU-instruction throughput probes (the epilogues of the tcgen05 kernels are one such instruction per accumulator column)
#define BNNB_MB_KERNEL(NAME, ASM)                                                                   \
  __global__ void __launch_bounds__(512) NAME(uint32_t* out, long long* cyc) {                      \
    uint32_t a[kUnroll], b = threadIdx.x | 1, c = blockIdx.x * 7 + 3;                               \
    _Pragma("unroll") for (int i = 0; i < kUnroll; i++) a[i] = threadIdx.x * 2654435761u + i;       \
    const long long t0 = clock64();                                                                 \
    for (int it = 0; it < kIters; it++) {                                                           \
      _Pragma("unroll") for (int i = 0; i < kUnroll; i++) asm volatile(ASM : "+r"(a[i]) : "r"(b), "r"(c)); \
    }                                                                                               \
    const long long t1 = clock64();                                                                 \
    uint32_t s = 0;                                                                                 \
    _Pragma("unroll") for (int i = 0; i < kUnroll; i++) s += a[i];                                  \
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;                                                 \
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;                                                \
  }
BNNB_MB_KERNEL(mb_shf, "shf.l.wrap.b32 %0, %1, %0, 1;")          // the sign-bit harvest: SHF.L.W.U32.HI
BNNB_MB_KERNEL(mb_prmt, "prmt.b32 %0, %0, %1, 0x7351;")
BNNB_MB_KERNEL(mb_imadhi, "mad.hi.u32 %0, %0, %1, %2;")
BNNB_MB_KERNEL(mb_imad, "mad.lo.u32 %0, %0, %1, %2;")
BNNB_MB_KERNEL(mb_shr, "shr.u32 %0, %0, 1;")
BNNB_MB_KERNEL(mb_fmnmx, "min.f32 %0, %0, %1;")
}  // namespace

extern "C" double bnnb_microbench(int which, double* clock_mhz) {
  cudaDeviceProp prop;
  int dev = 0;  // the calling thread's current device (a per-GPU process under torchrun has its own one current: never switch it)
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return -1;
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
    else if (which == 3) mb_shf<<<blocks, threads>>>(out, cyc);
    else if (which == 4) mb_prmt<<<blocks, threads>>>(out, cyc);
    else if (which == 5) mb_imadhi<<<blocks, threads>>>(out, cyc);
    else if (which == 6) mb_imad<<<blocks, threads>>>(out, cyc);
    else if (which == 7) mb_shr<<<blocks, threads>>>(out, cyc);
    else if (which == 8) mb_fmnmx<<<blocks, threads>>>(out, cyc);
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

// ---- TMEM read / write throughput: `warps` warps (1..8; warp w owns lane quarter w % 4) each issue `iters` tcgen05.ld (or .st)
// .32x32b.x32 (4 KB per warp-instruction) back to back.  Returns bytes per clock per SM (one CTA). --------------------------------
#include "tcgen05.cuh"
namespace {
__global__ void __launch_bounds__(256) k_tmem_rate(int warps, int iters, int store, long long* cyc, uint32_t* sink) {
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) bnnb::tc::tmem_alloc(&tmem_base, 512);
  bnnb::tc::fence_before_sync();
  __syncthreads();
  bnnb::tc::fence_after_sync();
  const uint32_t t0 = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 256;
  uint32_t r[32];
#pragma unroll
  for (int i = 0; i < 32; i++) r[i] = threadIdx.x + i;
  if (warp < warps) { bnnb::tc::tmem_st32(t0, r); bnnb::tc::tmem_st32(t0 + 32, r); bnnb::tc::tmem_st_wait(); }
  __syncthreads();
  const long long c0 = clock64();
  uint32_t acc = 0;
  if (warp < warps) {
    for (int it = 0; it < iters; it++) {
      if (store) {
        bnnb::tc::tmem_st32(t0 + (it & 7) * 32, r);
      } else {
        bnnb::tc::tmem_ld32(t0 + (it & 7) * 32, r);
      }
    }
    if (store) bnnb::tc::tmem_st_wait();
    else bnnb::tc::tmem_ld_wait_dep(r);
#pragma unroll
    for (int i = 0; i < 32; i++) acc ^= r[i];
  }
  __syncthreads();
  const long long c1 = clock64();
  if (threadIdx.x == 0) *cyc = c1 - c0;
  if (acc == 0x12345678u) sink[0] = acc;
  bnnb::tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) bnnb::tc::tmem_dealloc(tmem_base, 512);
}
}  // namespace

extern "C" double bnnb_tmem_rate(int warps, int store) {
  long long* cyc; uint32_t* sink;
  if (cudaMalloc(&cyc, 8) != cudaSuccess) return -1;
  cudaMalloc(&sink, 4);
  const int iters = 4096;
  k_tmem_rate<<<1, 256>>>(warps, iters, store, cyc, sink);
  k_tmem_rate<<<1, 256>>>(warps, iters, store, cyc, sink);
  long long c = 0;
  const cudaError_t e = cudaDeviceSynchronize();
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  cudaFree(cyc); cudaFree(sink);
  if (e != cudaSuccess || c <= 0) return -1;
  return (double)warps * iters * 4096.0 / (double)c;
}

// ---- mbarrier ping-pong: warp 0 arrives on A and waits on B, warp 1 waits on A and arrives on B.  Returns clocks per round trip
// (two hops).  mode 1: warp 0's arrival is a tcgen05.commit (no MMA outstanding) instead of a plain arrive. ---------------------------
namespace {
__global__ void __launch_bounds__(64) k_mbar_pingpong(int iters, int mode, long long* cyc) {
  __shared__ __align__(8) uint64_t bar[2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { bnnb::tc::mbar_init(&bar[0], 1); bnnb::tc::mbar_init(&bar[1], 1); bnnb::tc::mbar_init_fence(); }
  __syncthreads();
  const long long c0 = clock64();
  for (int i = 0; i < iters; i++) {
    if (warp == 0) {
      if (lane == 0) { if (mode == 1) bnnb::tc::commit(&bar[0]); else bnnb::tc::mbar_arrive(&bar[0]); }
      __syncwarp();
      bnnb::tc::mbar_wait(&bar[1], i & 1);
    } else {
      bnnb::tc::mbar_wait(&bar[0], i & 1);
      if (lane == 0) bnnb::tc::mbar_arrive(&bar[1]);
      __syncwarp();
    }
  }
  if (threadIdx.x == 0) *cyc = clock64() - c0;
}
}  // namespace

extern "C" double bnnb_mbar_roundtrip(int mode) {
  long long* cyc;
  if (cudaMalloc(&cyc, 8) != cudaSuccess) return -1;
  const int iters = 20000;
  k_mbar_pingpong<<<1, 64>>>(iters, mode, cyc);
  k_mbar_pingpong<<<1, 64>>>(iters, mode, cyc);
  long long c = 0;
  const cudaError_t e = cudaDeviceSynchronize();
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  cudaFree(cyc);
  return e == cudaSuccess ? (double)c / iters : -1;
}

// ---- tcgen05.ld .pack::16b probe: one warp writes in[lane][0..31] to 32 TMEM columns (tcgen05.st .x32) and reads them back with
// .32x32b.x16.pack::16b into out[lane][0..15].  Pins which 16 bits of which column land in which half of each register. -----------
namespace {
__global__ void __launch_bounds__(32) k_tmem_pack_probe(const uint32_t* in, uint32_t* out) {
  __shared__ uint32_t tmem_base;
  bnnb::tc::tmem_alloc(&tmem_base, 32);
  bnnb::tc::fence_before_sync();
  __syncwarp();
  bnnb::tc::fence_after_sync();
  uint32_t r[32], q[16];
#pragma unroll
  for (int i = 0; i < 32; i++) r[i] = in[threadIdx.x * 32 + i];
  bnnb::tc::tmem_st32(tmem_base, r);
  bnnb::tc::tmem_st_wait();
  bnnb::tc::tmem_ld32_pack16(tmem_base, q);
  bnnb::tc::tmem_ld_wait_dep16(q);
#pragma unroll
  for (int i = 0; i < 16; i++) out[threadIdx.x * 16 + i] = q[i];
  bnnb::tc::fence_before_sync();
  __syncwarp();
  bnnb::tc::tmem_dealloc(tmem_base, 32);
}
}  // namespace

extern "C" int bnnb_tmem_pack_probe(const uint32_t* in, uint32_t* out) {
  uint32_t *di, *dout;
  if (cudaMalloc(&di, 32 * 32 * 4) != cudaSuccess) return -1;
  cudaMalloc(&dout, 32 * 16 * 4);
  cudaMemcpy(di, in, 32 * 32 * 4, cudaMemcpyHostToDevice);
  k_tmem_pack_probe<<<1, 32>>>(di, dout);
  const bool ok = cudaDeviceSynchronize() == cudaSuccess;
  cudaMemcpy(out, dout, 32 * 16 * 4, cudaMemcpyDeviceToHost);
  cudaFree(di); cudaFree(dout);
  return ok ? 0 : -1;
}
