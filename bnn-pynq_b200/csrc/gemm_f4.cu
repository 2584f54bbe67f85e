// Fully-connected layers (and the 3x3 -> 1x1 convolution, which is one) on tcgen05.mma kind::mxf4 with unit scales: the exact
// small-integer MMA of conv_f4.cu, K-streamed.  Rows = images (tiles of 128), columns = output channels in blocks of NB.
//
// Warp-specialised persistent CTA (384 threads), decoupled by mbarriers only:
//   warps 0-3  epilogue: TMEM (thresholded layers: packed 16-bit reads of accumulators that carry the threshold in a biased pre-load,
//              f4_common.cuh) -> e2m1 codes (ACT_F4T) / decision bits / 16-bit scores -> global
//   warps 4-9  operand tile i+1 -> planes [K/32][128 rows][16 B], handed to the issuer in four K quarters so that the MMAs of a tile
//              start while its tail is still arriving: four TMA bulk copies when the producer wrote e2m1 tiles (ACT_F4T; one
//              thread), else expansion of bit-packed activations
//   warp  10   loader: weight slabs (NB x 64 e2m1 = NB*32 bytes, pre-arranged in operand layout by k_repack_w4s) by TMA bulk
//              copy into a ring of R slots, running ahead of the MMAs (the warp walks the groups in ring order, converged)
//   warp  11   issuer: one K=64 MMA per slab; accumulator slots (3) let the next channel block start while one drains
// Replaces StreamingFCLayer_Batch / ConvLayer_Batch + ThresholdsActivation / PassThroughActivation for these layers
// (fclayer.h:83-111, convlayer.h:85-121, activations.hpp:85-148).
#include <climits>
#include "kernels.cuh"
#include "tcgen05.cuh"
#include "f4_common.cuh"

namespace bnnb {
namespace {

#ifndef BNNB_GG
#define BNNB_GG 8
#endif
// Weight slabs per ring slot: one bulk copy, one full / empty hand-off and one tcgen05.commit per GROUP of slabs.  A wait + commit
// pair costs the issuer thread ~300 clocks while an MMA of these shapes runs 48-64: with one slab per slot the kernel with its MMAs
// and epilogue bodies compiled out took as long as the full one (lfcW1A1, 1024 x 1024 layer: 0.117 vs 0.128 ms; groups of 2 / 4 /
// 8 slabs: 0.086 / 0.071 / 0.058 ms).  The ring takes whatever shared memory is left (bytes in flight set the rate of the stream).
constexpr int GG = BNNB_GG;
// experiment switches (results wrong, hand-offs kept): -DBNNB_GEMM_STRIP=<bits> 1: no MMAs, 2: no accumulator reads / harvest / stores,
// 4: no pre-load, 8: no weight copies, 16: no operand expansion
#ifndef BNNB_GEMM_STRIP
#define BNNB_GEMM_STRIP 0
#endif
constexpr int kGStrip = BNNB_GEMM_STRIP;
// test-first mbarrier waits (tcgen05.cuh) by role: 1 loaders, 2 issuer, 4 epilogue
#ifndef BNNB_GEMM_TFM
#define BNNB_GEMM_TFM 0
#endif
constexpr bool gTF_L = (BNNB_GEMM_TFM & 1) != 0, gTF_I = (BNNB_GEMM_TFM & 2) != 0, gTF_E = (BNNB_GEMM_TFM & 4) != 0;
constexpr int GF_RING_MAX = 32;

// 8 bits -> 8 nibbles (bit i at position 4i)
__device__ __forceinline__ uint32_t gspread8(uint32_t x) {
  x = (x | (x << 12)) & 0x000f000fu;
  x = (x | (x << 6)) & 0x03030303u;
  x = (x | (x << 3)) & 0x11111111u;
  return x;
}
// 32 channels -> 16 bytes of e2m1.  IN_XNOR: bit -> {0, +1}; IN_BIN: bit -> {-1, +1}; IN_TERN: (nz, pos) -> {-1, 0, +1}
template <int INM>
__device__ __forceinline__ uint4 gexpand(uint32_t w0, uint32_t w1) {
  uint32_t v[4];
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const uint32_t a = gspread8((w0 >> (8 * q)) & 0xff);
    if (INM == IN_TERN) v[q] = (a << 1) | (gspread8(((w0 & ~w1) >> (8 * q)) & 0xff) << 3);
    else if (INM == IN_BIN) v[q] = 0xAAAAAAAAu ^ (a << 3);  // 1 -> 0x2, 0 -> 0xA
    else v[q] = a << 1;
  }
  return make_uint4(v[0], v[1], v[2], v[3]);
}

struct GemmGeom {
  int K, N, kch, ks, nblk, nbuf;  // kch = K/32 chunks, ks = K/64 slabs per channel block, nbuf = operand buffers (1 or 2)
  int cw, ipl;                    // words per pixel and plane (C/32), input planes
  int a_bytes, off_ring, off_thr, ring;  // ring = slots of the weight ring (GG slabs each)
  int resident;                          // the whole layer's weights fit the ring: loaded once per CTA, every tile re-uses them
  int* classes;                          // OUT_SCORE: optional class per image (first score strictly above a running max from 0,
  int ncls;                              // over the first ncls scores: foldedmv-offload.h:398-406), written by the thread that holds the row
};

template <int NB, int INM, int OUTM, bool TAP>
__global__ void __launch_bounds__(384, 1) k_gemm_f4(DevLayer L, GemmGeom g, const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int n_img,
                                                     int in_fmt, int out_fmt, int* __restrict__ status, int32_t* __restrict__ tap) {
  tc::griddep_launch();  // dependents (programmatic dependent launch) may start their prologues now
  constexpr int SLAB = NB * 32, SLOTS = 3, SF_COL = 384, NW = NB / 32, EXP = 192, U = 4;  // EXP expander threads, U loads in flight each
  // thresholded outputs: the accumulators are pre-loaded with the (first) threshold, biased so that bit 15 of the result's low half is
  // the decision; the epilogue reads two channels per register (.pack::16b) and forms codes / bit words with PRMT (f4_common.cuh)
  constexpr bool TERN = OUTM == OUT_TERN;
  constexpr bool ACC = OUTM != OUT_SCORE;  // the MMAs accumulate onto a pre-load
  constexpr int OPL = OUTM == OUT_TERN ? 2 : 1;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[SLOTS], empty[SLOTS], a_part[2][4], a_empty[2], ring_full[GF_RING_MAX], ring_empty[GF_RING_MAX];
  __shared__ uint32_t tmem_base;
  uint8_t* ring = smem + g.off_ring;
  float* sthr = reinterpret_cast<float*>(smem + g.off_thr);  // [2][N]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int role = warp < 4 ? 0 : (warp < 10 ? 1 : warp - 8);  // 0 epilogue, 1 expander, 2 loader, 3 issuer
  const int hw = warp & 3;

  for (int i = tid; i < 2 * g.N; i += 384) {
    const int t = L.thr_tc[i];
    float tf = t == INT_MAX ? 3.0e38f : (float)t;
    if (OUTM == OUT_SCORE) tf = (float)t;
    const int smax = 2 * g.K;  // |S| <= K * max|w| * max|a|
    if (!ACC) sthr[(i & 1) * g.N + (i >> 1)] = tf;
    else if (i < g.N) sthr[i] = tern_preload(L.thr_tc[2 * i], smax);                 // [0, N): pre-loads
    else if (TERN && i < g.N + g.N / 2) {                                                    // [N, N + N/2): (thr0 - thr1) of channels 2j, 2j+1
      const int j = i - g.N;
      reinterpret_cast<uint32_t*>(sthr)[i] = tern_delta(L.thr_tc[4 * j], L.thr_tc[4 * j + 1], L.thr_tc[4 * j + 2], L.thr_tc[4 * j + 3], smax);
    }
  }
  if (warp == 0) tc::tmem_alloc(&tmem_base, 512);
  if (tid == 0) {
    for (int i = 0; i < 2; i++) {
      tc::mbar_init(&a_empty[i], 1);
      for (int q = 0; q < 4; q++) tc::mbar_init(&a_part[i][q], in_fmt == ACT_F4T ? 1 : EXP);
    }
    for (int i = 0; i < SLOTS; i++) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 4); }
    for (int i = 0; i < g.ring; i++) { tc::mbar_init(&ring_full[i], 1); tc::mbar_init(&ring_empty[i], 1); }
    tc::mbar_init_fence();
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_base;
  if (warp < 4) tc::tmem_fill16(tmem + ((uint32_t)(warp * 32) << 16) + SF_COL, 0x7f7f7f7fu);  // every scale factor = 1.0 (ue8m0)
  const uint32_t sthr_u = tc::smem_u32(sthr);
  auto preload = [&](int slot, int nb) {  // this warp's lane quarter of accumulator `slot` <- -(thr + 1/2) of channel block nb
    const uint32_t t0 = tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * NB;
#pragma unroll
    for (int j = 0; j < NW; j++) {
      uint32_t v[32];
#pragma unroll
      for (int q = 0; q < 8; q++)
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(v[4 * q]), "=r"(v[4 * q + 1]), "=r"(v[4 * q + 2]), "=r"(v[4 * q + 3])
                     : "r"(sthr_u + (uint32_t)(nb * NB + j * 32 + q * 4) * 4));
      tc::tmem_st32(t0 + j * 32, v);
    }
    tc::tmem_st_wait();
  };
  const int ntiles = (n_img + 127) / 128;
  const int T = (int)blockIdx.x < ntiles ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;  // tiles of this CTA
  // pre-loads: the accumulator slot of the j-th channel block overall (j = tile_local * nblk + nb) holds that block's pre-load
  if (ACC && warp < 4)
    for (int j = 0; j < SLOTS && j < T * g.nblk; j++) preload(j, j % g.nblk);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t idesc = tc::make_idesc_mxf4(128, NB);
  const uint32_t sa_u0 = tc::smem_u32(smem), ring_u = tc::smem_u32(ring), sf = tmem + SF_COL;
  bool failed = false;

  if (role == 1) {
    // ---- expander: tile tl -> operand buffer tl % nbuf ------------------------------------------------------------------------
    const int et = tid - 128;  // 0..191
    tc::griddep_wait();  // the input is the kernel below's output (programmatic dependent launch)
    if (in_fmt == ACT_F4T) {
      // the producer wrote this tile's operand planes already (ACT_F4T): four TMA bulk copies, one per K quarter
      if (et == 0) {
        const uint8_t* in4 = reinterpret_cast<const uint8_t*>(in);
        for (int tl = 0; tl < T; tl++) {
          const int buf = tl % g.nbuf, use = tl / g.nbuf;
          if (use >= 1 && !failed && !tc::mbar_wait<gTF_L>(&a_empty[buf], (uint32_t)(use - 1) & 1)) failed = true;
          uint8_t* sa = smem + (size_t)buf * g.a_bytes;
          const uint8_t* src = in4 + ((size_t)blockIdx.x + (size_t)tl * gridDim.x) * g.a_bytes;
          for (int q = 0; q < 4; q++) {
            const int s0 = (q * g.ks) / 4, s1 = ((q + 1) * g.ks) / 4;
            if (s1 > s0) {
              tc::mbar_expect_tx(&a_part[buf][q], (uint32_t)(s1 - s0) * 4096);
              tc::bulk_g2s(sa + (size_t)s0 * 4096, src + (size_t)s0 * 4096, (uint32_t)(s1 - s0) * 4096, &a_part[buf][q]);
            } else {
              tc::mbar_arrive(&a_part[buf][q]);
            }
          }
        }
      }
    } else {
    const size_t img_words = (size_t)g.kch * g.ipl;
    const int total = 128 * g.ks;  // items: (row, pair of chunks = one K=64 slab), row fastest: conflict-free 16-byte stores
    for (int tl = 0; tl < T; tl++) {
      const int buf = tl % g.nbuf, use = tl / g.nbuf;
      if (use >= 1 && !failed && !tc::mbar_wait<gTF_L>(&a_empty[buf], (uint32_t)(use - 1) & 1)) failed = true;
      uint8_t* sa = smem + (size_t)buf * g.a_bytes;
      const long long img0 = ((long long)blockIdx.x + (long long)tl * gridDim.x) * 128;
      int qn = 0;
      for (int base = et; qn < 4; base += EXP * U) {
        uint2 w0[U], w1[U];
#pragma unroll
        for (int u = 0; u < U; u++) {  // U independent 8-byte loads in flight
          const int i = base + u * EXP, row = i & 127, j = (i >> 7) * 2;
          w0[u] = make_uint2(0, 0);
          w1[u] = make_uint2(0, 0);
          if (i < total && img0 + row < n_img) {
            const int p = j / g.cw, c = j - p * g.cw;
            const uint32_t* src = in + (size_t)(img0 + row) * img_words + (size_t)p * g.cw * g.ipl + c;
            w0[u] = __ldg(reinterpret_cast<const uint2*>(src));
            if (INM == IN_TERN) w1[u] = __ldg(reinterpret_cast<const uint2*>(src + g.cw));
          }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
          const int i = base + u * EXP, row = i & 127, j = (i >> 7) * 2;
          if (i < total) {
            *reinterpret_cast<uint4*>(sa + (size_t)j * 2048 + row * 16) = gexpand<INM>(w0[u].x, w1[u].x);
            *reinterpret_cast<uint4*>(sa + (size_t)(j + 1) * 2048 + row * 16) = gexpand<INM>(w0[u].y, w1[u].y);
          }
        }
        // this thread's next item is base + EXP*U: every quarter that ends at or before it is complete as far as this thread goes
        while (qn < 4 && base + EXP * U >= 128 * (((qn + 1) * g.ks) / 4)) {
          tc::fence_async_smem();
          tc::mbar_arrive(&a_part[buf][qn]);
          qn++;
        }
      }
    }
    }
  } else if (role == 2) {
    // ---- loader: weight slabs in MMA order, R slots ahead ----------------------------------------------------------------------
    // The slots come free in ring order, so the warp walks the groups in order, CONVERGED (every lane polls the same barrier; one
    // elected lane arms it and issues the copy).  With one lane per slot the lanes diverged, each spinning on its own barrier, and a
    // freed slot waited for its lane's turn.
    {
      const bool leader = tc::elect_one();
      const int ngrp = (g.ks + GG - 1) / GG;      // groups per channel block (the last one may be short)
      const int total = (g.resident ? (T > 0 ? 1 : 0) : T) * g.nblk * ngrp;  // resident weights: the first tile's groups fill the ring for good
      int r = 0, nb = 0, gi = 0;
      uint32_t rpar = 1;                           // parity to wait for on ring_empty = parity of the slot's previous use
      for (int gg = 0; gg < total; gg++) {
        const int s0 = gi * GG, cnt = min(GG, g.ks - s0);
        if (gg >= g.ring && !failed && !tc::mbar_wait<gTF_L>(&ring_empty[r], rpar)) failed = true;
        if (leader) {
          if (kGStrip & 8) tc::mbar_arrive(&ring_full[r]);
          else {
            tc::mbar_expect_tx(&ring_full[r], (uint32_t)cnt * SLAB);
            tc::bulk_g2s(ring + (size_t)r * (GG * SLAB), L.w4s + (size_t)(nb * g.ks + s0) * SLAB, (uint32_t)cnt * SLAB, &ring_full[r]);  // the slabs of a block are consecutive
          }
        }
        if (++gi == ngrp) { gi = 0; if (++nb == g.nblk) nb = 0; }
        if (++r == g.ring) { r = 0; rpar ^= 1u; }
      }
    }
  } else if (role == 3) {
    // ---- issuer: converged warp, one elected lane issues; one add per descriptor -----------------------------------------------
    {
      const bool leader = tc::elect_one();
      const uint32_t d_hi = tc::desc_hi(128);
      const uint32_t b_lo = tc::desc_lo(ring_u, NB * 16);
      int r = 0;
      uint32_t rpar = 0;  // ring slot and the parity of its current use: counted, not divided
      for (int tl = 0; tl < T; tl++) {
        const int buf = tl % g.nbuf, use = tl / g.nbuf;
        const uint32_t a_lo = tc::desc_lo(sa_u0 + (uint32_t)buf * g.a_bytes, 2048);
        for (int nb = 0; nb < g.nblk; nb++) {
          const int gb = tl * g.nblk + nb, slot = gb % SLOTS, u = gb / SLOTS;
          if (u >= 1 && !failed && !tc::mbar_wait<gTF_I>(&empty[slot], (uint32_t)(u - 1) & 1)) failed = true;
          tc::fence_after_sync();
          const uint32_t acc = tmem + (uint32_t)slot * NB;
          int qn = 0;
          int s_begin = 0;
#ifndef BNNB_GEMM_NO_LEAN
          // every channel block but a tile's first (all operand quarters are in): the full groups in a loop stripped to wait, one asm
          // block of eight MMAs, commit - the issuing thread retires an instruction every 6-7 clocks, and the general loop below
          // spends more of them per group than the eight MMAs take to run
          if (GG == 8 && nb > 0) {
            const int nfull = g.ks / GG;
            uint32_t a = a_lo;
            for (int gi = 0; gi < nfull; gi++) {
              if (!(g.resident && tl > 0) && !failed && !tc::mbar_wait<gTF_I>(&ring_full[r], rpar)) failed = true;
              if (leader) {
                if (!(kGStrip & 1)) tc::mma8_mxf4<256, NB * 2>(acc, a, b_lo + (uint32_t)r * (GG * (SLAB / 16)), d_hi, idesc, sf, (ACC || gi > 0) ? 1u : 0u);
                if (!g.resident) tc::commit(&ring_empty[r]);
              }
              a += 2 * GG * 128;
              if (++r == g.ring) { r = 0; rpar ^= 1u; }
            }
            __syncwarp();
            s_begin = nfull * GG;
            qn = 4;
          }
#endif
          for (int s0 = s_begin; s0 < g.ks; s0 += GG) {  // one ring slot = GG slabs (fewer at the end of the block)
            if (!(g.resident && tl > 0) && !failed && !tc::mbar_wait<gTF_I>(&ring_full[r], rpar)) failed = true;
            const int s1 = min(s0 + GG, g.ks);
#ifndef BNNB_GEMM_SLABWISE
            // The group as asm blocks of 8 / 4 / 2 MMAs (slab by slab with run-time indices the issuing thread spends ~260 clocks per
            // MMA, fc_chain.cu trace).  A tile's first channel block first waits for the operand quarters the group reads.
            if (GG == 8) {
              while (nb == 0 && qn < 4 && s1 - 1 >= (qn * g.ks) / 4) {
                if (!failed && !tc::mbar_wait<gTF_I>(&a_part[buf][qn], (uint32_t)use & 1)) failed = true;
                tc::fence_after_sync();
                qn++;
              }
              if (leader && !(kGStrip & 1)) {
                int e = 0;
                const int cnt = s1 - s0;
                const uint32_t a0 = a_lo + (uint32_t)(2 * s0) * 128, b0 = b_lo + (uint32_t)(r * GG) * (SLAB / 16);
                if (cnt - e >= 8) { tc::mma8_mxf4<256, NB * 2>(acc, a0, b0, d_hi, idesc, sf, (ACC || s0 > 0) ? 1u : 0u); e += 8; }
                if (cnt - e >= 4) { tc::mma4_mxf4<256, NB * 2>(acc, a0 + (uint32_t)(2 * e) * 128, b0 + (uint32_t)e * (SLAB / 16), d_hi, idesc, sf, (ACC || s0 + e > 0) ? 1u : 0u); e += 4; }
                if (cnt - e >= 2) { tc::mma2_mxf4<256, NB * 2>(acc, a0 + (uint32_t)(2 * e) * 128, b0 + (uint32_t)e * (SLAB / 16), d_hi, idesc, sf, (ACC || s0 + e > 0) ? 1u : 0u); e += 2; }
                if (cnt - e >= 1) tc::mma_mxf4_lh(acc, a0 + (uint32_t)(2 * e) * 128, d_hi, b0 + (uint32_t)e * (SLAB / 16), d_hi, idesc, sf, sf, (ACC || s0 + e > 0) ? 1u : 0u);
              }
            } else
#endif
            {
#ifndef BNNB_GEMM_NO_MMA8
            // a full group whose operand quarters are all in (every channel block but a tile's first): one asm block of eight MMAs
            if (GG % 8 == 0 && s1 - s0 == GG && (nb > 0 || qn == 4)) {
              if (leader && !(kGStrip & 1)) {
#pragma unroll
                for (int e = 0; e < GG / 8; e++)
                  tc::mma8_mxf4<256, NB * 2>(acc, a_lo + (uint32_t)(2 * (s0 + 8 * e)) * 128, b_lo + (uint32_t)(r * GG + 8 * e) * (SLAB / 16), d_hi, idesc, sf,
                                             (ACC || s0 + e > 0) ? 1u : 0u);
              }
            } else
#endif
            for (int s = s0; s < s1; s++) {
              // slab s lies in quarter q with ((q+1)*ks)/4 > s: wait for the quarters up to it (first channel block of a tile only)
              while (nb == 0 && qn < 4 && (s >= (qn * g.ks) / 4)) {
                if (!failed && !tc::mbar_wait<gTF_I>(&a_part[buf][qn], (uint32_t)use & 1)) failed = true;
                tc::fence_after_sync();
                qn++;
              }
              if (leader && !(kGStrip & 1))
                tc::mma_mxf4_lh(acc, a_lo + (uint32_t)(2 * s) * 128, d_hi, b_lo + (uint32_t)(r * GG + (s - s0)) * (SLAB / 16), d_hi, idesc, sf, sf, (ACC || s > 0) ? 1u : 0u);
            }
            }
            if (leader && !g.resident) tc::commit(&ring_empty[r]);
            if (++r == g.ring) { r = 0; rpar ^= 1u; }
            __syncwarp();
          }
          if (leader) tc::commit(&full[slot]);
          __syncwarp();
        }
        if (leader) tc::commit(&a_empty[buf]);
        __syncwarp();
      }
    }
  } else {
    // ---- epilogue: one image (row) per thread, every channel block ---------------------------------------------------------------
    const int nwt = g.N / 32;
    for (int tl = 0; tl < T; tl++) {
      const long long img = ((long long)blockIdx.x + (long long)tl * gridDim.x) * 128 + hw * 32 + lane;
      int cls_max = 0, cls_arg = 0;  // OUT_SCORE with classes: running arg-max over this row's scores, channel order
      for (int nb = 0; nb < g.nblk; nb++) {
        const int gb = tl * g.nblk + nb, slot = gb % SLOTS, u = gb / SLOTS;
        if (!failed && !tc::mbar_wait<gTF_E>(&full[slot], (uint32_t)u & 1)) failed = true;
        tc::fence_after_sync();
        const uint32_t trow = tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * NB;
        uint32_t b0[NW], b1[NW];
        uint4 c4[NW];  // e2m1 codes instead of bits when the consumer is another gemm_f4 launch
        const bool codes = ACC && out_fmt == ACT_F4T;
        uint32_t r[2][32];  // scores, TAP
        uint32_t hh[2][16];
        if (kGStrip & 2) {}
        else if (ACC) tc::tmem_ld32_pack16(trow, hh[0]);
        else tc::tmem_ld32(trow, r[0]);
#pragma unroll
        for (int j = 0; j < NW; j++) {
          const float* th = sthr + nb * NB + j * 32;
          uint32_t w0 = 0, w1 = 0;
          c4[j] = make_uint4(0, 0, 0, 0);
          if (kGStrip & 2) { b0[j] = b1[j] = 0; continue; }
          if (ACC) {
            uint32_t dl[16];  // 2-bit outputs: the deltas of this block and word, from shared memory while the load is in flight
            if (TERN) {
              const uint4* dp = reinterpret_cast<const uint4*>(reinterpret_cast<const uint32_t*>(sthr) + g.N + (nb * NB + j * 32) / 2);
#pragma unroll
              for (int q = 0; q < 4; q++) { const uint4 v = dp[q]; dl[4 * q] = v.x; dl[4 * q + 1] = v.y; dl[4 * q + 2] = v.z; dl[4 * q + 3] = v.w; }
            }
            tc::tmem_ld_wait_dep16(hh[j & 1]);
            if (j + 1 < NW) tc::tmem_ld32_pack16(trow + (j + 1) * 32, hh[(j + 1) & 1]);
            if (TERN) {
              if (codes) c4[j] = tern_codes16(hh[j & 1], dl);
              else tern_planes16(hh[j & 1], dl, w0, w1);
            } else {
              if (codes) c4[j] = bin_codes16(hh[j & 1]);
              else w0 = tern_bit_word(hh[j & 1]);
            }
            if (TAP) { tc::tmem_ld32(trow + j * 32, r[j & 1]); tc::tmem_ld_wait_dep(r[j & 1]); }
          } else {
            tc::tmem_ld_wait_dep(r[j & 1]);
            if (j + 1 < NW) tc::tmem_ld32(trow + (j + 1) * 32, r[(j + 1) & 1]);  // in flight while word j is formed
            if (img < n_img) {  // PassThroughActivation<ap_uint<16>>: (S + offset) mod 2^16, offset = #negative weights for 1-bit inputs
              int16_t* o16 = reinterpret_cast<int16_t*>(out) + (size_t)img * g.N + nb * NB + j * 32;
#pragma unroll
              for (int c = 0; c < 32; c += 8) {
                uint32_t pk[4];
#pragma unroll
                for (int e = 0; e < 4; e++) {
                  const uint32_t lo = (uint32_t)(__float2int_rn(__uint_as_float(r[j & 1][c + 2 * e])) + __float2int_rn(th[c + 2 * e])) & 0xffff;
                  const uint32_t hi = (uint32_t)(__float2int_rn(__uint_as_float(r[j & 1][c + 2 * e + 1])) + __float2int_rn(th[c + 2 * e + 1])) & 0xffff;
                  pk[e] = lo | (hi << 16);
                  const int ch = nb * NB + j * 32 + c + 2 * e, vlo = (int)(int16_t)lo, vhi = (int)(int16_t)hi;
                  if (ch < g.ncls && vlo > cls_max) { cls_max = vlo; cls_arg = ch; }
                  if (ch + 1 < g.ncls && vhi > cls_max) { cls_max = vhi; cls_arg = ch + 1; }
                }
                *reinterpret_cast<uint4*>(o16 + c) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              }
            }
          }
          b0[j] = w0;
          b1[j] = w1;
          if (TAP && img < n_img) {  // parity taps (TAP instantiation only): the MMA sum of every channel, pre-load un-folded
            int32_t* ta = tap + (size_t)img * g.N + nb * NB + j * 32;
#pragma unroll
            for (int c = 0; c < 32; c++) ta[c] = __float2int_rn(__uint_as_float(r[j & 1][c]) - (ACC ? th[c] : 0.0f));
          }
        }
        if (OUTM == OUT_SCORE && g.classes && nb == g.nblk - 1 && img < n_img) g.classes[img] = cls_arg;
        if (ACC && !(kGStrip & 4) && gb + SLOTS < T * g.nblk) preload(slot, (gb + SLOTS) % g.nblk);
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&empty[slot]);  // this warp's quarter of the accumulator has been read (and re-loaded)
        if (codes && img < n_img) {  // ACT_F4T: [tile][chunk = channel / 32][row]
          uint4* out4 = reinterpret_cast<uint4*>(out) + (((size_t)blockIdx.x + (size_t)tl * gridDim.x) * nwt + (size_t)nb * NW) * 128 + hw * 32 + lane;
#pragma unroll
          for (int j = 0; j < NW; j++) out4[(size_t)j * 128] = c4[j];
        } else if (OUTM != OUT_SCORE && img < n_img) {
          if (OUTM == OUT_BIT && NW == 2 && g.classes && g.nblk == 1) {
            // LFC: the class of the 64-bit output word, multi-image rule (foldedmv-offload.cpp:202-220; same arithmetic as k_lfc_classes)
            const uint64_t w = (((uint64_t)b0[1] << 32) | b0[0]) & (g.ncls >= 64 ? ~0ull : (g.ncls <= 0 ? 0ull : (~0ull >> (64 - g.ncls))));
            g.classes[img] = !w ? 0 : (w < (1ull << 32) ? 31 - __clz((unsigned)w) : (int)(unsigned)log2((double)w));
          }
          uint32_t* dst = out + (size_t)img * nwt * OPL + nb * NW;
#pragma unroll
          for (int j = 0; j < NW; j += 2) {
            *reinterpret_cast<uint2*>(dst + j) = make_uint2(b0[j], b0[j + 1]);
            if (OUTM == OUT_TERN) *reinterpret_cast<uint2*>(dst + nwt + j) = make_uint2(b1[j], b1[j + 1]);
          }
        }
      }
    }
  }
  if (failed && status) *status = 1;
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

template <int NB, int INM, int OUTM>
bool launch_g(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int in_fmt, int out_fmt, int* status, int32_t* tap,
              cudaStream_t s, int* classes, int ncls) {
  const LayerDesc& d = L.d;
  GemmGeom g;
  g.classes = (OUTM == OUT_SCORE || (OUTM == OUT_BIT && d.N == 64)) ? classes : nullptr;  // CNV scores / the LFC output word
  g.ncls = ncls < 0 ? 0 : (ncls > d.N ? d.N : ncls);
  g.K = d.K; g.N = d.N; g.kch = d.K / 32; g.ks = d.K / 64; g.nblk = d.N / NB;
  g.cw = d.ifm_ch / 32; g.ipl = INM == IN_TERN ? 2 : 1;
  g.a_bytes = g.kch * 2048;
  const int thr = 2 * d.N * 4, budget = 224 * 1024, slot_bytes = GG * NB * 32, ring_min = 2 * slot_bytes;
  // two operand buffers only if four ring slots remain beside them: the weight stream needs bytes in flight more than the operand
  // tile needs a second buffer (measured round 2: two slots + two buffers 0.063 ms against four slots + one buffer 0.056 ms per
  // 1024 x 1024 layer; slots of four slabs instead of eight: 0.072 ms)
  g.nbuf = 2 * g.a_bytes + 2 * ring_min + thr <= budget ? 2 : 1;
  if (g.nbuf * g.a_bytes + ring_min + thr > budget) return false;
  g.ring = (budget - g.nbuf * g.a_bytes - thr) / slot_bytes;
  if (g.ring > GF_RING_MAX) g.ring = GF_RING_MAX;
  // small layers (CNV layers 6-8: 16-128 KB of weights): the ring holds every group of the layer, loaded once per CTA instead of
  // once per tile (measured: 0.023 -> 0.021 ms per 512 x 512 layer - these kernels are mostly launch and set-up; a second epilogue
  // team on the idle expander warps changed nothing either)
  const int groups = g.nblk * ((g.ks + GG - 1) / GG);
  g.resident = groups <= g.ring ? 1 : 0;
  if (g.resident) g.ring = groups;
  const int ring = g.ring * slot_bytes;
  g.off_ring = g.nbuf * g.a_bytes;
  g.off_thr = g.off_ring + ring;
  const int smem = g.off_thr + thr;
  const int ntiles = (n_img + 127) / 128;
  const int grid = ntiles < num_sms ? ntiles : num_sms;
  if (tap) {  // parity taps: the TAP instantiation of the same kernel
    auto kern = k_gemm_f4<NB, INM, OUTM, true>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    kern<<<grid, 384, smem, s>>>(L, g, in, out, n_img, in_fmt, out_fmt, status, tap);
    return true;
  }
  auto kern = k_gemm_f4<NB, INM, OUTM, false>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  launch_maybe_pdl(kern, grid, 384, (size_t)smem, s, L, g, in, out, n_img, in_fmt, out_fmt, status, (int32_t*)nullptr);
  return true;
}

template <int NB>
bool launch_nb(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int in_fmt, int out_fmt, int* status, int32_t* tap,
               cudaStream_t s, int* classes, int ncls) {
  const int im = L.d.in_mode, om = L.d.out_mode;
  if (im == IN_XNOR && om == OUT_BIT) return launch_g<NB, IN_XNOR, OUT_BIT>(L, in, out, n_img, num_sms, in_fmt, out_fmt, status, tap, s, classes, ncls);
  if (im == IN_XNOR && om == OUT_SCORE) return launch_g<NB, IN_XNOR, OUT_SCORE>(L, in, out, n_img, num_sms, in_fmt, out_fmt, status, tap, s, classes, ncls);
  if (im == IN_BIN && om == OUT_TERN) return launch_g<NB, IN_BIN, OUT_TERN>(L, in, out, n_img, num_sms, in_fmt, out_fmt, status, tap, s, classes, ncls);
  if (im == IN_TERN && om == OUT_TERN) return launch_g<NB, IN_TERN, OUT_TERN>(L, in, out, n_img, num_sms, in_fmt, out_fmt, status, tap, s, classes, ncls);
  if (im == IN_TERN && om == OUT_SCORE) return launch_g<NB, IN_TERN, OUT_SCORE>(L, in, out, n_img, num_sms, in_fmt, out_fmt, status, tap, s, classes, ncls);
  if (im == IN_TERN && om == OUT_BIT) return launch_g<NB, IN_TERN, OUT_BIT>(L, in, out, n_img, num_sms, in_fmt, out_fmt, status, tap, s, classes, ncls);
  return false;
}

}  // namespace

int gemm_f4_block(const LayerDesc& d) { return d.N % 128 == 0 ? 128 : 64; }

// Returns false when the layer cannot run on this kernel (the caller falls back to the next core).
bool gemm_f4_supports(const DevLayer& L) {
  const LayerDesc& d = L.d;
  if (!L.w4s || !L.thr_tc || d.pool || d.ofm_dim != 1 || d.K % 64 || d.ifm_ch % 64 || d.N % 64 || L.kpad != d.K) return false;
  if (d.K != d.ifm_ch * d.ifm_dim * d.ifm_dim) return false;  // the im2col row must be the whole activation tensor of the image
  const int im = d.in_mode, om = d.out_mode;
  return (im == IN_XNOR && (om == OUT_BIT || om == OUT_SCORE)) || (im == IN_BIN && om == OUT_TERN) || (im == IN_TERN);
}

bool launch_gemm_f4(const DevLayer& L, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int in_fmt, int out_fmt, int* status, int32_t* tap,
                    cudaStream_t s, int* classes, int ncls) {
  const LayerDesc& d = L.d;
  if (!gemm_f4_supports(L) || n_img <= 0) return false;
  if (in_fmt == ACT_F4P || out_fmt == ACT_F4P) return false;
  if (in_fmt == ACT_F4T && d.in_mode != IN_XNOR && d.in_mode != IN_TERN) return false;  // {0,1} / {-1,0,+1} codes as the producer wrote them
  if (out_fmt == ACT_F4T && d.out_mode != OUT_BIT && d.out_mode != OUT_TERN) return false;
  return gemm_f4_block(d) == 128 ? launch_nb<128>(L, in, out, n_img, num_sms, in_fmt, out_fmt, status, tap, s, classes, ncls) : launch_nb<64>(L, in, out, n_img, num_sms, in_fmt, out_fmt, status, tap, s, classes, ncls);
}

}  // namespace bnnb
