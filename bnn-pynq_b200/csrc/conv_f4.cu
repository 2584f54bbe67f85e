// conv3x3 layers on tcgen05.mma kind::mxf4 (block-scaled fp4 with every scale = 1.0): an exact small-integer MMA at twice the
// int8 rate.  Activations {0,1} / {-1,0,+1} and weights {-1,+1} / {-2..+1} are all exactly representable in e2m1; products and
// the fp32 accumulation are exact (|acc| <= 2*2304 << 2^24), so the thresholded outputs are bit-identical to the reference's
// integer MVAU (mvau.hpp:86-172, activations.hpp:139-147).  Evidence: tests/test_gpu_tcgen05.py::test_fp4_probe_*.
//
// Same implicit GEMM as conv_tc.cu (operand planes of 16 bytes = 32 channels per pixel, tap (ky,kx) = descriptor start shifted by
// (ky*IDIM+kx) pixels), now one K=64 MMA per 64 input channels per tap, and a warp-specialised CTA (one per SM, 640 threads):
//   warps 0-15      epilogue: groups of four warps (one per TMEM lane quarter), either four channel quarters on every tile or two
//                   channel halves x two teams on alternate tiles (a lone warp retires an instruction every 6-7 clocks, so the
//                   per-tile epilogue is split until it is shorter than the tile's MMAs): TMEM (packed 16-bit reads of accumulators
//                   that carry the threshold in a biased pre-load, f4_common.cuh) -> e2m1 codes or decision bits -> global
//                   (pooled layers: via smem)
//   warps 16-18     operand planes of group g+1 into the other buffer, the group's images interleaved pixel by pixel: one tiled TMA
//                   load per plane through a tensor map (ACT_F4P producer) or one bulk copy per plane (ACT_F4I producer), issued by
//                   one thread; else expansion of bit-packed activations (global, L2) -> fp4 operand planes
//   warp  19        MMA issuer, up to three tiles ahead (three TMEM accumulator slots)
// decoupled by mbarriers only: abuf_full / abuf_empty per operand buffer, full / empty per accumulator slot.
// PAIR instantiation: two CTAs of a cluster run every MMA together (tcgen05 cta_group::2, M = 256), see the kernel.
#include <cuda.h>  // CUtensorMap (the type and the encoder's signature only: the entry point comes from cudaGetDriverEntryPoint)
#include <climits>
#include <cstdlib>
#include "kernels.cuh"
#include "tcgen05.cuh"
#include "f4_common.cuh"

namespace bnnb {
namespace {

// The G images of a group are interleaved pixel by pixel: operand row = pixel * G + image (ACT_F4I input: one bulk copy per plane;
// ACT_F4P input: one tiled TMA load per plane through a tensor map; bit input: expanded into that order).  The rows whose outputs
// can be valid - pixel index up to (OD-1)*IDIM + OD - then form a prefix of OROWS rows, and only its tiles are computed (the bottom
// rows of every image were 14-45 % of these layers' MMAs); tap (ky, kx) = descriptor start (ky * IDIM + kx) * G rows on.
template <int C, int N, int NS, int IDIM, int OD, bool POOL, int G, int INM, int OUTM>
struct F4Cfg {
  static constexpr int PXI = IDIM * IDIM, ROWS = G * PXI, XS = G, YS = IDIM * XS, MAXSHIFT = 2 * YS + 2 * XS;
  static constexpr int OROWS = ((OD - 1) * IDIM + OD) * G;  // rows 0 .. OROWS-1 hold every valid output
  static constexpr int TILES = (OROWS + 127) / 128;
  static constexpr int PITCH = ((TILES * 128 + MAXSHIFT > ROWS ? TILES * 128 + MAXSHIFT : ROWS) + 7) / 8 * 8;
  static constexpr int KC = C / 32;  // 16-byte chunks (32 fp4 channels) per tap
  static constexpr int NT = N * NS;  // output channels of the layer: NS blocks of N (one MMA accumulator each)
  static constexpr int PLANE = PITCH * 16, A_BYTES = KC * PLANE, B_BYTES = 9 * KC * NT * 16;
  static constexpr int IPL = INM == IN_TERN ? 2 : 1, OPL = OUTM == OUT_TERN ? 2 : 1, NW = N / 32, NWT = NT / 32, WPP = NWT * OPL;
  static constexpr int BITS_BYTES = POOL ? TILES * 128 * WPP * 4 : 0;  // staging rows (pooled layers only)
  static constexpr int OFF_B = 2 * A_BYTES, OFF_BITS = OFF_B + B_BYTES, OFF_THR = OFF_BITS + BITS_BYTES, SMEM = OFF_THR + 2 * NT * 4;
  static constexpr int SLOTS = 384 / N > 4 ? 4 : 384 / N;  // accumulator slots below the scale-factor columns
  static constexpr int SF_COL = 384, PD = POOL ? OD / 2 : OD;
};

// experiment switches (results wrong, hand-offs kept): -DBNNB_CONV_STRIP=<bits> 1: no MMAs, 2: no TMEM loads + harvest, 4: no pre-load,
// 8: no global stores, 16: no input copies, 32: no pooling (its barriers kept), 64: no pooling barriers either
#ifndef BNNB_CONV_STRIP
#define BNNB_CONV_STRIP 0
#endif
constexpr int kStrip = BNNB_CONV_STRIP;
// test-first mbarrier waits (tcgen05.cuh) by role: 1 loader, 2 issuer, 4 epilogue
#ifndef BNNB_CONV_TFM
#define BNNB_CONV_TFM 7
#endif
constexpr bool kTF_L = (BNNB_CONV_TFM & 1) != 0, kTF_I = (BNNB_CONV_TFM & 2) != 0, kTF_E = (BNNB_CONV_TFM & 4) != 0;
#ifndef BNNB_G2B
#define BNNB_G2B 10
#endif
#ifndef BNNB_G2T
#define BNNB_G2T 10
#endif
#ifndef BNNB_G4
#define BNNB_G4 19
#endif
#ifndef BNNB_G3T
#define BNNB_G3T 6
#endif
#ifndef BNNB_CONV_EGS
#define BNNB_CONV_EGS 0
#endif
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

template <int C, int N, int NS, int IDIM, int OD, bool POOL, int G, int INM, int OUTM, bool TAP, bool PAIR>
__global__ void __launch_bounds__(640, 1) k_conv_f4(DevLayer L, const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int n_img,
                                                     int in_fmt, int out_fmt, int og, int* __restrict__ status, int32_t* __restrict__ tap,
                                                     const __grid_constant__ CUtensorMap tmap) {
  tc::griddep_launch();  // dependents (programmatic dependent launch) may start their prologues now
  using Cfg = F4Cfg<C, N, NS, IDIM, OD, POOL, G, INM, OUTM>;
  const uint32_t og_magic = f4i_magic(og > 0 ? og : 1);  // ACT_F4I output: the consumer's images per group
  static_assert(Cfg::NW % 4 == 0, "pooled rows of decision bits are moved as uint4");
  static_assert(BNNB_CONV_EGS == 4 || BNNB_CONV_EGS == 2 || BNNB_CONV_EGS == 0, "epilogue groups");
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[4], empty[4], abuf_full[2], abuf_empty[2], peer_full[2];
  __shared__ uint32_t tmem_base;
  uint8_t* sb = smem + Cfg::OFF_B;
  float* sthr = reinterpret_cast<float*>(smem + Cfg::OFF_THR);  // [2][N]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int role = warp < 16 ? 0 : (warp < 19 ? 1 : 2), ht = tid & 127, hw = warp & 3;
  // epilogue warps (role 0): EGS channel groups x TEAMS tile teams of four warps (one per TMEM lane quarter); team t takes the tiles
  // with gt % TEAMS == t, group eg the output channels [eg*N/EGS, (eg+1)*N/EGS) of the block
  // (measured with the interleaved layout, ms per 65 536 images 4 x 1 -> 2 x 2: layer 2 0.312 -> 0.305 (1-bit), 0.393 -> 0.374 (2-bit) - the
  // fixed per-tile cost of a warp is halved; the pooled layer 3 0.375 -> 0.386, 0.375 -> 0.436 - its tiles are twice as long already)
  constexpr int EGS = BNNB_CONV_EGS ? BNNB_CONV_EGS : (POOL ? 4 : (NS == 2 ? (OUTM == OUT_BIT ? 2 : 4) : 2)), TEAMS = 4 / EGS;
  const int eg = ((warp >> 2) & 3) % EGS, team = ((warp >> 2) & 3) / EGS;
  constexpr int NWG = Cfg::NW / EGS;  // 32-channel words per epilogue group
  constexpr int NG = N / EGS;         // channels per epilogue group
  constexpr int KCT = 9 * Cfg::KC;
  // The accumulators are pre-loaded with the (first) threshold, biased so that the low 16 bits of the result hold the decision
  // thr < S in bit 15 (f4_common.cuh), and every MMA accumulates.  The epilogue reads two channels per register (.pack::16b: the
  // TMEM read port, 64 B per clock and SM, bounds a 128 x 128 tile read as 32-bit words at 1024 clocks against 576 of MMAs) and
  // turns sign bits into codes / bit words with PRMT; 2-bit outputs get the second threshold with one packed add.
  constexpr bool TERN = OUTM == OUT_TERN;
  constexpr int SMAX = 2 * 9 * C;        // |S| <= K * max|w| * max|a|

  constexpr int NT = Cfg::NT;
  // PAIR (cta_group::2, ACT_F4P inputs): the two CTAs of a cluster - the two SMs of a TPC - run every MMA together with M = 256: each
  // keeps its own image group (operand rows, accumulator lanes) and HALF of every channel block's weight rows, so an MMA costs each
  // SM 32 + 32 shared-memory wavefronts instead of 32 + 64 (these shapes are operand-fetch-bound: 64 per 64 clocks of math at N = 128).
  // The leader (rank 0) issues and commits to both CTAs' barriers; the follower's loader and epilogue warps arrive on the leader's.
  const uint32_t rank = PAIR ? tc::cluster_ctarank() : 0;
  constexpr int NTL = PAIR ? NT / 2 : NT, NL = PAIR ? N / 2 : N;  // weight rows this CTA holds: per layer, per channel block
  for (int i = tid; i < NTL * KCT; i += 640) {  // weights (packed e2m1, [N][K/2]) -> chunk-major operand layout
    const int nl = i % NTL, kc = i / NTL;
    const int n = PAIR ? (nl / NL) * N + (int)rank * NL + nl % NL : nl;
    *reinterpret_cast<uint4*>(sb + ((size_t)kc * NTL + nl) * 16) = *reinterpret_cast<const uint4*>(L.w4 + (size_t)n * (L.kpad / 2) + kc * 16);
  }
  for (int i = tid; i < 2 * NT; i += 640) {
    if (i < NT) sthr[i] = tern_preload(L.thr_tc[2 * i], SMAX);                        // [0, NT): pre-loads
    else if (TERN && i < NT + NT / 2) {                                                        // [NT, NT + NT/2): (thr0 - thr1) of channels 2j, 2j+1
      const int j = i - NT;
      reinterpret_cast<uint32_t*>(sthr)[i] = tern_delta(L.thr_tc[4 * j], L.thr_tc[4 * j + 1], L.thr_tc[4 * j + 2], L.thr_tc[4 * j + 3], SMAX);
    }
  }
  for (int i = tid; i < 2 * Cfg::A_BYTES / 16; i += 640) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);  // pad pixels stay 0
  if (PAIR) tc::cluster_sync();  // both CTAs are resident before the paired allocation
  if (warp == 0) { if (PAIR) tc::tmem_alloc_pair(&tmem_base, 512); else tc::tmem_alloc(&tmem_base, 512); }
  if (tid == 0) {
    for (int i = 0; i < 2; i++) { tc::mbar_init(&abuf_full[i], (in_fmt == ACT_F4I || in_fmt == ACT_F4PM) ? 1 : 96); tc::mbar_init(&abuf_empty[i], 1); tc::mbar_init(&peer_full[i], 1); }
    for (int i = 0; i < 4; i++) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], (PAIR ? 2 : 1) * 4 * EGS); }
    tc::mbar_init_fence();
  }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  if (PAIR) tc::cluster_sync();  // the peer's barriers are initialised and its weights staged before anything crosses
  tc::fence_after_sync();
  const uint32_t tmem = tmem_base;
  if (warp < 4) tc::tmem_fill16(tmem + ((uint32_t)(warp * 32) << 16) + Cfg::SF_COL, 0x7f7f7f7fu);  // every scale factor = 1.0 (ue8m0)
  const uint32_t sthr_u = tc::smem_u32(sthr);
  // One 32-channel word per warp and one channel block: the warp's 32 pre-load values never change, so they live in registers.  (A
  // shared-memory read waits behind the MMAs' operand stream: re-reading them per tile cost the fused kernel +50 %.)
#ifdef BNNB_CONV_PRE_SMEM
  constexpr bool PRE_REG = false;
#else
  constexpr bool PRE_REG = NWG == 1 && NS == 1;
#endif
  uint32_t pre[32];
  if (PRE_REG) {
#pragma unroll
    for (int c = 0; c < 32; c++) pre[c] = __float_as_uint(sthr[(warp < 16 ? eg : 0) * NG + c]);
  }
  auto preload = [&](int slot, int h) {  // this warp's lane quarter and channel half of accumulator `slot` <- -(thr + 1/2) of block h
    const uint32_t t0 = tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * N;
    if (PRE_REG) {
      tc::tmem_st32(t0 + eg * 32, pre);
      tc::tmem_st_wait();
      return;
    }
#pragma unroll
    for (int jj = 0; jj < NWG; jj++) {
      const int j = eg * NWG + jj;
      uint32_t v[32];
#pragma unroll
      for (int q = 0; q < 8; q++) {
        // volatile: keeps the 128 loop-invariant values out of registers
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(v[4 * q]), "=r"(v[4 * q + 1]), "=r"(v[4 * q + 2]), "=r"(v[4 * q + 3])
                     : "r"(sthr_u + (uint32_t)(h * N + j * 32 + q * 4) * 4));
      }
      tc::tmem_st32(t0 + j * 32, v);
    }
    tc::tmem_st_wait();
  };
  uint32_t dl[16];  // 2-bit outputs, one word per warp and one channel block: the deltas live in registers too
  if (TERN && PRE_REG) {
#pragma unroll
    for (int c = 0; c < 16; c++) dl[c] = reinterpret_cast<const uint32_t*>(sthr)[NT + ((warp < 16 ? eg : 0) * NG) / 2 + c];
  }
  if (warp < 16)
    for (int slot = 0; slot < Cfg::SLOTS; slot++)
      if (slot % TEAMS == team) preload(slot, slot % NS);
  tc::fence_before_sync();
  __syncthreads();
  if (PAIR) tc::cluster_sync();  // the peer's pre-loads are in place before the leader's first MMA accumulates onto them
  tc::fence_after_sync();
  const uint32_t idesc = tc::make_idesc_mxf4(PAIR ? 256 : 128, N);
  const uint32_t sa_u0 = tc::smem_u32(smem), sb_u = tc::smem_u32(sb), sf = tmem + Cfg::SF_COL;
  bool failed = false;
  const int ngroups = (n_img + G - 1) / G;
  // groups of this CTA: k-th group = gfirst + k * gstride (PAIR: the cluster takes pairs of groups, rank r the r-th of each pair; both
  // CTAs run the same number of rounds - a group past the end is empty: nothing loaded, nothing stored)
  const int nunits = PAIR ? (ngroups + 1) / 2 : ngroups, unit0 = PAIR ? (int)blockIdx.x >> 1 : (int)blockIdx.x, ustride = PAIR ? (int)gridDim.x >> 1 : (int)gridDim.x;
  const int K = unit0 < nunits ? (nunits - unit0 + ustride - 1) / ustride : 0;
  const int gfirst = PAIR ? 2 * unit0 + (int)rank : unit0, gstride = PAIR ? 2 * ustride : ustride;

  if (role == 1) {
    // ---- operand planes of group k+1, rows interleaved (pixel, image): ACT_F4P input = one tiled TMA load per plane through the tensor
    // map, ACT_F4I input = one bulk copy per plane (both issued by one thread); ACT_BITS input = three warps expanding bit-packed
    // activations --------------------------------------------------------------------------------------------------------------
    constexpr int CW = C / 32;
    tc::griddep_wait();  // the input is the kernel below's output (programmatic dependent launch)
    if (in_fmt == ACT_F4I || in_fmt == ACT_F4PM) {
      if (tid == 512) {
        const uint8_t* in4 = reinterpret_cast<const uint8_t*>(in);
        constexpr uint32_t IMG_PLANE = Cfg::PXI * 16;  // bytes of one plane of one image
        for (int k = 0; k < K; k++) {
          const int img0 = (gfirst + k * gstride) * G, buf = k & 1;
          if (k >= 2 && !failed && !tc::mbar_wait<kTF_L>(&abuf_empty[buf], (uint32_t)((k >> 1) - 1) & 1)) failed = true;
          uint8_t* sa = smem + (size_t)buf * Cfg::A_BYTES;
          const int valid = n_img - img0 < G ? (n_img - img0 > 0 ? n_img - img0 : 0) : G;  // rows of missing images keep stale (finite) codes; never stored
          if (in_fmt == ACT_F4PM) {
            // ACT_F4P tensor [img][plane][pixel][16 B] read through a 4-D tensor map with dimensions (16 B, image, pixel, plane): one
            // box of (16 B, G images, all pixels, one plane) lands in shared memory in box order = rows (pixel, image); images past
            // the end of the tensor are filled with zeros
            if (kStrip & 16) tc::mbar_arrive(&abuf_full[buf]); else tc::mbar_expect_tx(&abuf_full[buf], (uint32_t)CW * G * IMG_PLANE);
            for (int j = 0; j < CW; j++)
              if (!(kStrip & 16)) tc::tma_load_4d(sa + (size_t)j * Cfg::PLANE, &tmap, 0, img0, 0, j, &abuf_full[buf]);
          } else {  // ACT_F4I: the group's plane j is one contiguous block [pixel][image] (slots of missing images hold stale codes)
            const int cp = valid > 0 ? CW : 0;
            if (kStrip & 16) tc::mbar_arrive(&abuf_full[buf]); else tc::mbar_expect_tx(&abuf_full[buf], (uint32_t)cp * G * IMG_PLANE);
            for (int j = 0; j < cp; j++)
              if (!(kStrip & 16)) tc::bulk_g2s(sa + (size_t)j * Cfg::PLANE, in4 + ((size_t)(img0 / G) * CW + j) * (G * IMG_PLANE), G * IMG_PLANE, &abuf_full[buf]);
          }
          if (PAIR && rank != 0) {  // the follower tells the leader's issuer once its planes have landed
            if (!failed && !tc::mbar_wait<kTF_L>(&abuf_full[buf], (uint32_t)(k >> 1) & 1)) failed = true;
            tc::mbar_arrive_remote(&peer_full[buf], 0);
          }
        }
      }
    } else
    for (int k = 0; k < K; k++) {
      const int img0 = (gfirst + k * gstride) * G, buf = k & 1;
      if (k >= 2 && !failed && !tc::mbar_wait<kTF_L>(&abuf_empty[buf], (uint32_t)((k >> 1) - 1) & 1)) failed = true;
      uint8_t* sa = smem + (size_t)buf * Cfg::A_BYTES;
#pragma unroll 4
      for (int i = tid - 512; i < Cfg::ROWS * CW; i += 96) {
        const int j = i % CW, mi = i / CW, gi = mi / Cfg::PXI, px = mi - gi * Cfg::PXI;  // (image, pixel) in the order of the global tensor
        uint32_t w0 = 0, w1 = 0;
        if (img0 + gi < n_img) {
          const size_t base = ((size_t)(img0 + gi) * Cfg::PXI + px) * (CW * Cfg::IPL);
          w0 = __ldg(in + base + j);
          if (INM == IN_TERN) w1 = __ldg(in + base + CW + j);
        }
        const int m = px * G + gi;  // operand row
        *reinterpret_cast<uint4*>(sa + (size_t)j * Cfg::PLANE + (size_t)m * 16) = expand_fp4<INM>(w0, w1);
      }
      tc::fence_async_smem();
      tc::mbar_arrive(&abuf_full[buf]);
    }
  } else if (role == 2) {
    // ---- issuer: the whole warp runs the schedule (uniform control flow), one elected lane issues; descriptors as (lo, hi) words
    // so that the address arithmetic per MMA is one add (a lone thread issues about one instruction per five clocks) ------------
    if (!PAIR || rank == 0) {
      const bool leader = tc::elect_one();
      const uint32_t d_hi = tc::desc_hi(128);
      const uint32_t a_lo[2] = {tc::desc_lo(sa_u0, Cfg::PLANE), tc::desc_lo(sa_u0 + Cfg::A_BYTES, Cfg::PLANE)};
      const uint32_t b_lo = tc::desc_lo(sb_u, NTL * 16);
      for (int k = 0; k < K; k++) {
        const int buf = k & 1;
        if (!failed && !tc::mbar_wait<kTF_I>(&abuf_full[buf], (uint32_t)(k >> 1) & 1)) failed = true;
        if (PAIR && !failed && !tc::mbar_wait<kTF_I>(&peer_full[buf], (uint32_t)(k >> 1) & 1)) failed = true;
        for (int vt = 0; vt < Cfg::TILES * NS; vt++) {  // virtual tile = (row tile t, channel block h)
          const int t = vt / NS, h = vt % NS;
          const int gt = k * Cfg::TILES * NS + vt, slot = gt % Cfg::SLOTS, u = gt / Cfg::SLOTS;
          if (u >= 1 && !failed && !tc::mbar_wait<kTF_I>(&empty[slot], (uint32_t)(u - 1) & 1)) failed = true;
          tc::fence_after_sync();
          const uint32_t acc = tmem + (uint32_t)slot * N;
          const uint32_t at = a_lo[buf] + (uint32_t)(t * 128), bt = b_lo + (uint32_t)(h * NL);
          if (leader) {  // nine taps per asm block (tcgen05.cuh: mma9_mxf4), once per pair of operand planes
#pragma unroll
            for (int s = 0; s < Cfg::KC / 2; s++)
              if (!(kStrip & 1)) tc::mma9_mxf4<Cfg::XS, Cfg::YS, Cfg::KC * NTL, PAIR ? 2 : 1>(acc, at + (uint32_t)(2 * s * (Cfg::PLANE / 16)), bt + (uint32_t)(2 * s * NTL), d_hi, idesc, sf,
                                               1u);
            if (PAIR) tc::commit_pair(&full[slot]); else tc::commit(&full[slot]);
          }
          __syncwarp();
        }
        if (leader) { if (PAIR) tc::commit_pair(&abuf_empty[buf]); else tc::commit(&abuf_empty[buf]); }
        __syncwarp();
      }
    }
  } else {
    // ---- epilogue groups: every tile, channels [eg*N/2, (eg+1)*N/2) --------------------------------------------------------------
    uint32_t* sbits = reinterpret_cast<uint32_t*>(smem + Cfg::OFF_BITS);
    constexpr int NW = Cfg::NW, NWT = Cfg::NWT, WPP = Cfg::WPP;
    for (int k = 0; k < K; k++) {
      const int img0 = (gfirst + k * gstride) * G;
      for (int vt = 0; vt < Cfg::TILES * NS; vt++) {
        const int t = vt / NS, h = vt % NS;
        const int gt = k * Cfg::TILES * NS + vt, slot = gt % Cfg::SLOTS, u = gt / Cfg::SLOTS;
        if (TEAMS > 1 && gt % TEAMS != team) continue;
        if (!failed && !tc::mbar_wait<kTF_E>(&full[slot], (uint32_t)u & 1)) failed = true;
        tc::fence_after_sync();
        const int m = t * 128 + hw * 32 + lane;
        const uint32_t trow = tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * N + (uint32_t)eg * NG;
        uint32_t b0[NWG], b1[NWG];
        uint4 c4[NWG];  // e2m1 codes instead of bits when the consumer is another fp4 kernel
        const bool codes = !POOL && out_fmt != ACT_BITS;  // ACT_F4P / ACT_F4I / ACT_F4T
        uint32_t r[32];  // TAP only
        uint32_t hh[2][16];
        if (!(kStrip & 2)) tc::tmem_ld32_pack16(trow, hh[0]);
#pragma unroll
        for (int j = 0; j < NWG; j++) {
          uint32_t w0 = 0, w1 = 0;
          c4[j] = make_uint4(0, 0, 0, 0);
          if (kStrip & 2) { b0[j] = b1[j] = 0; continue; }
          if (TERN && !PRE_REG) {  // the deltas of this block and word, from shared memory while the load is in flight
            const uint4* dp = reinterpret_cast<const uint4*>(reinterpret_cast<const uint32_t*>(sthr) + NT + (h * N + eg * NG + j * 32) / 2);
#pragma unroll
            for (int q = 0; q < 4; q++) { const uint4 v = dp[q]; dl[4 * q] = v.x; dl[4 * q + 1] = v.y; dl[4 * q + 2] = v.z; dl[4 * q + 3] = v.w; }
          }
          tc::tmem_ld_wait_dep16(hh[j & 1]);
          if (j + 1 < NWG) tc::tmem_ld32_pack16(trow + (j + 1) * 32, hh[(j + 1) & 1]);  // in flight while word j is formed
          if (TERN) {
            if (codes) c4[j] = tern_codes16(hh[j & 1], dl);
            else tern_planes16(hh[j & 1], dl, w0, w1);
          } else {
            if (codes) c4[j] = bin_codes16(hh[j & 1]);
            else w0 = tern_bit_word(hh[j & 1]);
          }
          if (TAP) { tc::tmem_ld32(trow + j * 32, r); tc::tmem_ld_wait_dep(r); }
          b0[j] = w0;
          b1[j] = w1;
          if (TAP) {  // parity taps (TAP instantiation only): the MMA sum of every valid output pixel and channel, pre-load un-folded
            const int gi = m % G, rem = m / G, y = rem / IDIM, x = rem - y * IDIM;
            if (m < Cfg::OROWS && y < OD && x < OD && img0 + gi < n_img) {
              int32_t* ta = tap + ((size_t)(img0 + gi) * (OD * OD) + y * OD + x) * NT + h * N + eg * NG + j * 32;
              const float* pv = sthr + h * N + eg * NG + j * 32;
#pragma unroll
              for (int c = 0; c < 32; c++) ta[c] = __float2int_rn(__uint_as_float(r[c]) - pv[c]);
            }
          }
        }
        if (!(kStrip & 4)) preload(slot, (gt + Cfg::SLOTS) % NS);
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) { if (PAIR && rank != 0) tc::mbar_arrive_remote(&empty[slot], 0); else tc::mbar_arrive(&empty[slot]); }  // this warp's share has been read (and re-loaded)
        uint32_t* row = nullptr;
        if (POOL) {
          row = sbits + (size_t)m * WPP;
        } else {  // straight to global: one output pixel (half of its channels) per thread
          const int gi = m % G, rem = m / G, y = rem / IDIM, x = rem - y * IDIM;
          const bool valid = m < Cfg::OROWS && y < OD && x < OD && img0 + gi < n_img;
          if (valid && !codes && !(kStrip & 8)) row = out + ((size_t)(img0 + gi) * (OD * OD) + y * OD + x) * WPP;
          if (valid && codes && !(kStrip & 8)) {
            const size_t img = (size_t)(img0 + gi);
            const int px = y * OD + x;
            uint4* out4 = reinterpret_cast<uint4*>(out);
            // ACT_F4I: one division per thread and tile, 32-bit index arithmetic (the launcher refuses tensors of 2^31 rows or more)
            const uint32_t fgrp = __umulhi((uint32_t)img, og_magic), fil = (uint32_t)img - fgrp * (uint32_t)og;
            const uint32_t fbase = ((fgrp * NWT + (uint32_t)(h * NW + eg * NWG)) * (OD * OD) + (uint32_t)px) * (uint32_t)og + fil, fstep = (uint32_t)(OD * OD) * (uint32_t)og;
#pragma unroll
            for (int j = 0; j < NWG; j++) {
              const int pj = h * NW + eg * NWG + j;  // 32-channel plane of this layer's output
              if (out_fmt == ACT_F4P) out4[(img * NWT + pj) * (OD * OD) + px] = c4[j];
              else if (out_fmt == ACT_F4I) out4[fbase + (uint32_t)j * fstep] = c4[j];
              else out4[((img >> 7) * (OD * OD * NWT) + (size_t)(px * NWT + pj)) * 128 + (img & 127)] = c4[j];  // ACT_F4T
            }
          }
        }
        if (row) {
          if (NWG == 1) {
            row[h * NW + eg] = b0[0];
            if (OUTM == OUT_TERN) row[NWT + h * NW + eg] = b1[0];
          } else if (NWG == 2) {
            *reinterpret_cast<uint2*>(row + h * NW + eg * NWG) = make_uint2(b0[0], b0[1]);
            if (OUTM == OUT_TERN) *reinterpret_cast<uint2*>(row + NWT + h * NW + eg * NWG) = make_uint2(b1[0], b1[1]);
          } else {
#pragma unroll
            for (int q = 0; q < NWG / 4; q++) {
              reinterpret_cast<uint4*>(row + h * NW + eg * NWG)[q] = make_uint4(b0[4 * q], b0[4 * q + 1], b0[4 * q + 2], b0[4 * q + 3]);
              if (OUTM == OUT_TERN) reinterpret_cast<uint4*>(row + NWT + h * NW + eg * NWG)[q] = make_uint4(b1[4 * q], b1[4 * q + 1], b1[4 * q + 2], b1[4 * q + 3]);
            }
          }
        }
      }
      if (POOL && !(kStrip & 64)) {
        asm volatile("bar.sync 1, 512;" ::: "memory");
        constexpr int PP = Cfg::PD * Cfg::PD;
        for (int i = tid; i < G * PP && !(kStrip & 32); i += 512) {  // one pooled output pixel per thread
          const int gi = i / PP, op = i - gi * PP, oy = op / Cfg::PD, ox = op - oy * Cfg::PD;
          if (img0 + gi >= n_img) continue;
          const uint4* r00 = reinterpret_cast<const uint4*>(sbits + (size_t)(((2 * oy) * IDIM + 2 * ox) * G + gi) * WPP);
          const uint4 *r01 = r00 + (size_t)Cfg::XS * (WPP / 4), *r10 = r00 + (size_t)Cfg::YS * (WPP / 4), *r11 = r10 + (size_t)Cfg::XS * (WPP / 4);
          uint4* dst = reinterpret_cast<uint4*>(out + ((size_t)(img0 + gi) * PP + op) * WPP);
          auto or4 = [](uint4 a, uint4 b) { return make_uint4(a.x | b.x, a.y | b.y, a.z | b.z, a.w | b.w); };
          auto and4 = [](uint4 a, uint4 b) { return make_uint4(a.x & b.x, a.y & b.y, a.z & b.z, a.w & b.w); };
          auto andn4 = [](uint4 a, uint4 b) { return make_uint4(a.x & ~b.x, a.y & ~b.y, a.z & ~b.z, a.w & ~b.w); };
#pragma unroll
          for (int q = 0; q < NWT / 4; q++) {
            if (OUTM == OUT_BIT) {
              const uint4 o = or4(or4(r00[q], r01[q]), or4(r10[q], r11[q]));
              if (out_fmt == ACT_F4P) {  // [img][plane][pixel], planes 4q .. 4q+3
                uint4* out4 = reinterpret_cast<uint4*>(out) + ((size_t)(img0 + gi) * NWT + 4 * q) * PP + op;
                out4[0] = f4_from_bits(o.x); out4[PP] = f4_from_bits(o.y); out4[2 * PP] = f4_from_bits(o.z); out4[3 * PP] = f4_from_bits(o.w);
              } else if (out_fmt == ACT_F4I) {  // [img / og][plane][pixel][img % og]
                uint4* out4 = reinterpret_cast<uint4*>(out) + f4i_index((uint32_t)(img0 + gi), 4 * q, op, NWT, PP, og, og_magic);
                const size_t ps = (size_t)PP * og;
                out4[0] = f4_from_bits(o.x); out4[ps] = f4_from_bits(o.y); out4[2 * ps] = f4_from_bits(o.z); out4[3 * ps] = f4_from_bits(o.w);
              } else {
                dst[q] = o;
              }
            } else {  // max over {-1,0,+1}: +1 if any +1; else -1 only if all four are -1
              constexpr int H = NWT / 4;
              const uint4 P = or4(or4(r00[H + q], r01[H + q]), or4(r10[H + q], r11[H + q]));
              const uint4 M = and4(and4(andn4(r00[q], r00[H + q]), andn4(r01[q], r01[H + q])), and4(andn4(r10[q], r10[H + q]), andn4(r11[q], r11[H + q])));
              if (out_fmt == ACT_F4P || out_fmt == ACT_F4I) {  // planes 4q .. 4q+3, as e2m1 codes
                const size_t ps = out_fmt == ACT_F4P ? (size_t)PP : (size_t)PP * og;
                uint4* out4 = reinterpret_cast<uint4*>(out) +
                              (out_fmt == ACT_F4P ? ((size_t)(img0 + gi) * NWT + 4 * q) * PP + op : f4i_index((uint32_t)(img0 + gi), 4 * q, op, NWT, PP, og, og_magic));
                out4[0] = expand_fp4<IN_TERN>(P.x | M.x, P.x); out4[ps] = expand_fp4<IN_TERN>(P.y | M.y, P.y);
                out4[2 * ps] = expand_fp4<IN_TERN>(P.z | M.z, P.z); out4[3 * ps] = expand_fp4<IN_TERN>(P.w | M.w, P.w);
              } else {
                dst[q] = or4(P, M);
                dst[H + q] = P;
              }
            }
          }
        }
        asm volatile("bar.sync 1, 512;" ::: "memory");  // the next image group overwrites the staging rows
      }
    }
  }
  if (failed && status) *status = 1;
  tc::fence_before_sync();
  __syncthreads();
  if (PAIR) tc::cluster_sync();  // the leader's MMAs read this CTA's shared memory and write its accumulators until the very end
  if (warp == 0) { if (PAIR) tc::tmem_dealloc_pair(tmem, 512); else tc::tmem_dealloc(tmem, 512); }
}

// Tensor map over an ACT_F4P tensor uint4 [n_img][planes][px], dimensions (16 bytes, image, pixel, plane), box (16, g, px, 1): the
// box lands in shared memory with dimension 0 fastest, i.e. as operand rows (pixel, image).  cuTensorMapEncodeTiled is taken from
// the driver through the runtime (no link against libcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
bool encode_f4p_map(CUtensorMap* m, const void* base, int n_img, int planes, int px, int g) {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
    return (EncodeTiledFn)p;
  }();
  if (!fn) return false;
  const cuuint64_t dims[4] = {16, (cuuint64_t)n_img, (cuuint64_t)px, (cuuint64_t)planes};
  const cuuint64_t strides[3] = {(cuuint64_t)planes * px * 16, 16, (cuuint64_t)px * 16};  // bytes, dimensions 1 .. 3
  const cuuint32_t box[4] = {16, (cuuint32_t)g, (cuuint32_t)px, 1}, estr[4] = {1, 1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int C, int N, int NS, int IDIM, int OD, bool POOL, int G, int INM, int OUTM>
bool launch_cfg(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int in_fmt, int out_fmt, int og, int* status, int32_t* tap,
                cudaStream_t s) {
  using Cfg = F4Cfg<C, N, NS, IDIM, OD, POOL, G, INM, OUTM>;
  static_assert(Cfg::SMEM <= 226 * 1024, "shared memory budget");
  static_assert(Cfg::SLOTS * N <= Cfg::SF_COL, "TMEM budget");
  const int ngroups = (n_img + G - 1) / G;
  const int grid = ngroups < num_sms ? ngroups : num_sms;
  CUtensorMap tmap = {};
  if (in_fmt == ACT_F4PM && !encode_f4p_map(&tmap, in, n_img, C / 32, IDIM * IDIM, G)) return false;
  // CTA pairs where they pay.  Measured, ms per 65 536 images, single -> paired: 1-bit layers 2 / 3 / 4: 0.335 -> 0.334, 0.413 -> 0.425,
  // 0.166 -> 0.140; 2-bit: 0.461 -> 0.503, 0.460 -> 0.398, 0.191 -> 0.176.  BNN_B200_CONV_PAIR=<mask> overrides (bit 0 / 1 / 2: layer 2 / 3 / 4)
  static const int pair_env = getenv("BNN_B200_CONV_PAIR") ? atoi(getenv("BNN_B200_CONV_PAIR")) : -1;
  const int pair_mask = pair_env >= 0 ? pair_env : (OUTM == OUT_TERN ? 6 : 4), pair_bit = NS == 2 ? 4 : (POOL ? 2 : 1);
  if ((in_fmt == ACT_F4I || in_fmt == ACT_F4PM) && ngroups >= 2 && (pair_mask & pair_bit)) {  // a cluster of two per TPC, each pair takes two image groups per round
    const int npairs = (ngroups + 1) / 2, maxpairs = num_sms / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * (npairs < maxpairs ? npairs : maxpairs));
    cfg.blockDim = dim3(640);
    cfg.dynamicSmemBytes = Cfg::SMEM;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (tap) {
      auto kern = k_conv_f4<C, N, NS, IDIM, OD, POOL, G, INM, OUTM, true, true>;
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
      return cudaLaunchKernelEx(&cfg, kern, L, in, out, n_img, in_fmt, out_fmt, og, status, tap, tmap) == cudaSuccess;
    }
    auto kern = k_conv_f4<C, N, NS, IDIM, OD, POOL, G, INM, OUTM, false, true>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
    return cudaLaunchKernelEx(&cfg, kern, L, in, out, n_img, in_fmt, out_fmt, og, status, (int32_t*)nullptr, tmap) == cudaSuccess;
  }
  if (tap) {  // parity taps: the TAP instantiation of the same kernel
    auto kern = k_conv_f4<C, N, NS, IDIM, OD, POOL, G, INM, OUTM, true, false>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
    kern<<<grid, 640, Cfg::SMEM, s>>>(L, in, out, n_img, in_fmt, out_fmt, og, status, tap, tmap);
    return true;
  }
  auto kern = k_conv_f4<C, N, NS, IDIM, OD, POOL, G, INM, OUTM, false, false>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
  launch_maybe_pdl(kern, grid, 640, Cfg::SMEM, s, L, in, out, n_img, in_fmt, out_fmt, og, status, (int32_t*)nullptr, tmap);
  return true;
}

}  // namespace

bool conv_f4_supports(const LayerDesc& d) {
  if (d.kind != KIND_CONV) return false;
  const bool a1 = d.in_mode == IN_XNOR && d.out_mode == OUT_BIT, a2 = d.in_mode == IN_TERN && d.out_mode == OUT_TERN;
  if (!a1 && !a2) return false;
  return (d.ifm_ch == 64 && d.N == 128 && d.ifm_dim == 14 && !d.pool) || (d.ifm_ch == 128 && d.N == 128 && d.ifm_dim == 12 && d.pool) ||
         (d.ifm_ch == 128 && d.N == 256 && d.ifm_dim == 5 && !d.pool);
}

// images per group of the layer's kernel: what a producer needs to write ACT_F4I for it (0: not a conv_f4 layer)
int conv_f4_group(const LayerDesc& d) {
  if (!conv_f4_supports(d)) return 0;
  const bool tern = d.out_mode == OUT_TERN;
  if (d.ifm_dim == 14) return tern ? BNNB_G2T : BNNB_G2B;
  if (d.ifm_dim == 12) return tern ? BNNB_G3T : 6;
  return BNNB_G4;
}

bool launch_conv_f4(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int in_fmt, int out_fmt, int og, int* status, int32_t* tap,
                    cudaStream_t s) {
  const LayerDesc& d = L.d;
  if (!conv_f4_supports(d) || !L.w4 || !L.thr_tc || n_img <= 0) return false;
  if (in_fmt == ACT_F4T || ((d.pool) && out_fmt == ACT_F4T)) return false;
  if (in_fmt == ACT_F4P) in_fmt = ACT_F4PM;  // image planes are always staged through the tensor map (interleaved rows)
  if (out_fmt == ACT_F4I && (og <= 0 || (long long)(n_img + og) * (d.N / 32) * d.ofm_dim * d.ofm_dim >= (1ll << 31))) return false;
  const bool a1 = d.in_mode == IN_XNOR && d.out_mode == OUT_BIT, a2 = d.in_mode == IN_TERN && d.out_mode == OUT_TERN;
  if (d.ifm_ch == 64 && d.N == 128 && d.ifm_dim == 14 && !d.pool) {
    if (a1) return launch_cfg<64, 128, 1, 14, 12, false, BNNB_G2B, IN_XNOR, OUT_BIT>(L, in, out, n_img, num_sms, in_fmt, out_fmt, og, status, tap, s);
    if (a2) return launch_cfg<64, 128, 1, 14, 12, false, BNNB_G2T, IN_TERN, OUT_TERN>(L, in, out, n_img, num_sms, in_fmt, out_fmt, og, status, tap, s);
  }
  if (d.ifm_ch == 128 && d.N == 128 && d.ifm_dim == 12 && d.pool) {
    if (a1) return launch_cfg<128, 128, 1, 12, 10, true, 6, IN_XNOR, OUT_BIT>(L, in, out, n_img, num_sms, in_fmt, out_fmt, og, status, tap, s);
    if (a2) return launch_cfg<128, 128, 1, 12, 10, true, BNNB_G3T, IN_TERN, OUT_TERN>(L, in, out, n_img, num_sms, in_fmt, out_fmt, og, status, tap, s);
  }
  if (d.ifm_ch == 128 && d.N == 256 && d.ifm_dim == 5 && !d.pool) {  // two channel blocks of 128 share the operand planes
    if (a1) return launch_cfg<128, 128, 2, 5, 3, false, BNNB_G4, IN_XNOR, OUT_BIT>(L, in, out, n_img, num_sms, in_fmt, out_fmt, og, status, tap, s);
    if (a2) return launch_cfg<128, 128, 2, 5, 3, false, BNNB_G4, IN_TERN, OUT_TERN>(L, in, out, n_img, num_sms, in_fmt, out_fmt, og, status, tap, s);
  }
  return false;
}

}  // namespace bnnb
