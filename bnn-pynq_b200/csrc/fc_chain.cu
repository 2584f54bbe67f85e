// A run of fully-connected layers in ONE launch (CNV layers 6-8, LFC layers 1-3): the 128-image tile never leaves the SM between
// layers.  Same machinery as gemm_f4.cu - tcgen05.mma kind::mxf4 with unit scales, weights streamed through a ring of TMA bulk
// copies, accumulators pre-loaded with the biased threshold, packed 16-bit epilogue reads (f4_common.cuh) - but a hidden layer's
// epilogue writes its e2m1 codes straight into the NEXT layer's operand tile in shared memory ([K/32][128 rows][16 B], the layout
// ACT_F4T has in global memory), block by block: the next layer's first channel block starts on the K slabs that are already
// there, so the tensor pipe does not drain at a layer boundary.  One launch replaces three: each of them has a fill-and-drain
// floor of 10-15 us (weights and thresholds in, first tile through copy -> MMAs -> epilogue) that a 65 536-image chunk pays once
// per layer and a single image pays in full.
//
// Warp roles (256 threads): warps 0-3 epilogue (one image row per thread), warp 4 first operand tile (TMA, one thread), warp 5
// weight loader, warp 6 issuer.  Replaces the StreamingFCLayer_Batch chain of DoCompute (cnvW1A1/hw/top.cpp:228-232,
// lfcW1A1/hw/top.cpp:142-148; fclayer.h:83-111, activations.hpp:85-148).
#include <algorithm>
#include <climits>
#include <cstdio>
#include "kernels.cuh"
#include "tcgen05.cuh"
#include "f4_common.cuh"

namespace bnnb {
namespace {

#ifndef BNNB_CH_GG
#define BNNB_CH_GG 8
#endif
constexpr int CH_MAX = 4, CH_GG = BNNB_CH_GG, CH_SLOTS = 3, CH_SF_COL = 384, CH_RING_MAX = 16, CH_BLK_MAX = 32;
// test-first mbarrier waits (tcgen05.cuh) by role: 1 loaders, 2 issuer, 4 epilogue
#ifndef BNNB_CHAIN_TFM
#define BNNB_CHAIN_TFM 0
#endif
constexpr bool cTF_L = (BNNB_CHAIN_TFM & 1) != 0, cTF_I = (BNNB_CHAIN_TFM & 2) != 0, cTF_E = (BNNB_CHAIN_TFM & 4) != 0;
constexpr int CH_SLOT_BYTES = CH_GG * 128 * 32;  // one ring slot: eight K=64 slabs of 128 channels

struct ChainGeom {
  int nl, nblk_tile, nbuf0, rel_layer;  // rel_layer: the first operand buffer is free again after this layer's MMAs
  int K[CH_MAX], N[CH_MAX], ks[CH_MAX], nblk[CH_MAX], nb[CH_MAX], blk0[CH_MAX];
  int a_off[CH_MAX];    // layer l's operand tile (layer 0: buffer 0, buffer 1 follows it)
  int thr_off[CH_MAX];  // index of layer l's table in sthr: N pre-loads (or score offsets), then N/2 packed deltas (2-bit outputs)
  int a_bytes0, off_ring, off_thr, ring;
  const uint8_t* w4s[CH_MAX];
  const int32_t* thr_tc[CH_MAX];
  int* classes;
  int ncls;
};

// TERN: hidden layers write 2-bit activations; LASTM: OUT_SCORE (CNV: 16-bit scores) or OUT_BIT (LFC: the 64-bit output word)
template <bool TERN, int LASTM>
__global__ void __launch_bounds__(256, 1) k_fc_chain(ChainGeom g, const uint8_t* __restrict__ in, uint32_t* __restrict__ out, int n_img,
                                                      int* __restrict__ status) {
  tc::griddep_launch();  // dependents (programmatic dependent launch) may start their prologues now
  constexpr int SLOTS = CH_SLOTS;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[SLOTS], empty[SLOTS], a_part[2][4], a_empty[2], a_ready[CH_MAX][8], ring_full[CH_RING_MAX],
      ring_empty[CH_RING_MAX];
  __shared__ uint32_t tmem_base;
  __shared__ uint8_t blk_layer[CH_BLK_MAX], blk_nb[CH_BLK_MAX];
  uint8_t* ring = smem + g.off_ring;
  float* sthr = reinterpret_cast<float*>(smem + g.off_thr);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, hw = warp & 3;
  const int last = g.nl - 1;
  auto is_acc = [&](int l) { return l < last || LASTM != OUT_SCORE; };  // the MMAs accumulate onto a pre-load

  for (int l = 0; l < g.nl; l++) {
    const int N = g.N[l], smax = 2 * g.K[l];
    const int32_t* tt = g.thr_tc[l];
    float* st = sthr + g.thr_off[l];
    const bool tern = TERN && l < last;
    for (int i = tid; i < N + (tern ? N / 2 : 0); i += 256) {
      if (i < N) st[i] = is_acc(l) ? tern_preload(tt[2 * i], smax) : (float)tt[2 * i];
      else {
        const int j = i - N;
        reinterpret_cast<uint32_t*>(st)[i] = tern_delta(tt[4 * j], tt[4 * j + 1], tt[4 * j + 2], tt[4 * j + 3], smax);
      }
    }
  }
  if (warp == 0) tc::tmem_alloc(&tmem_base, 512);
  if (tid == 0) {
    for (int i = 0; i < 2; i++) {
      tc::mbar_init(&a_empty[i], 1);
      for (int q = 0; q < 4; q++) tc::mbar_init(&a_part[i][q], 1);
    }
    for (int i = 0; i < SLOTS; i++) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 4); }
    for (int i = 0; i < g.ring; i++) { tc::mbar_init(&ring_full[i], 1); tc::mbar_init(&ring_empty[i], 1); }
    for (int l = 0; l < CH_MAX; l++)
      for (int b = 0; b < 8; b++) tc::mbar_init(&a_ready[l][b], 128);
    for (int l = 0; l < g.nl; l++)
      for (int b = 0; b < g.nblk[l]; b++) { blk_layer[g.blk0[l] + b] = (uint8_t)l; blk_nb[g.blk0[l] + b] = (uint8_t)b; }
    tc::mbar_init_fence();
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_base;
  if (warp < 4) tc::tmem_fill16(tmem + ((uint32_t)(warp * 32) << 16) + CH_SF_COL, 0x7f7f7f7fu);  // every scale factor = 1.0 (ue8m0)
  const uint32_t sthr_u = tc::smem_u32(sthr);
  auto preload = [&](int slot, int l, int nb) {  // this warp's lane quarter of accumulator `slot` <- pre-loads of block nb of layer l
    const uint32_t t0 = tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * 128;
    const int nw = g.nb[l] / 32;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      if (j < nw) {
        uint32_t v[32];
#pragma unroll
        for (int q = 0; q < 8; q++)
          asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(v[4 * q]), "=r"(v[4 * q + 1]), "=r"(v[4 * q + 2]), "=r"(v[4 * q + 3])
                       : "r"(sthr_u + (uint32_t)(g.thr_off[l] + nb * g.nb[l] + j * 32 + q * 4) * 4));
        tc::tmem_st32(t0 + j * 32, v);
      }
    }
    tc::tmem_st_wait();
  };
  const int ntiles = (n_img + 127) / 128;
  const int T = (int)blockIdx.x < ntiles ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;  // tiles of this CTA
  const int nbt = g.nblk_tile;
  if (warp < 4)
    for (int j = 0; j < SLOTS && j < T * nbt; j++) {
      const int b = j % nbt;
      if (is_acc(blk_layer[b])) preload(j, blk_layer[b], blk_nb[b]);
    }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t sa_u0 = tc::smem_u32(smem), ring_u = tc::smem_u32(ring), sf = tmem + CH_SF_COL;
  bool failed = false;
#ifdef BNNB_CHAIN_TRACE
  __shared__ long long tr[3][16][6];  // [role: 0 issuer, 1 epilogue, 2 loader][block of the traced tile][event]
  constexpr int TRT = 1;              // the traced tile of CTA 0
#define CTR(role, b, e) do { if (blockIdx.x == 0 && tl == TRT && lane == 0 && (b) < 16) tr[role][b][e] = clock64(); } while (0)
#else
#define CTR(role, b, e) do {} while (0)
#endif

  if (warp == 4) {
    // ---- first operand tile of every tile of this CTA: four TMA bulk copies, one per K quarter (the producer wrote ACT_F4T) -------
    if (lane == 0) {
      tc::griddep_wait();  // the input is the kernel below's output (programmatic dependent launch)
      const int ks = g.ks[0];
      for (int tl = 0; tl < T; tl++) {
        const int buf = tl % g.nbuf0, use = tl / g.nbuf0;
        if (use >= 1 && !failed && !tc::mbar_wait<cTF_L>(&a_empty[buf], (uint32_t)(use - 1) & 1)) failed = true;
        uint8_t* sa = smem + g.a_off[0] + (size_t)buf * g.a_bytes0;
        const uint8_t* src = in + ((size_t)blockIdx.x + (size_t)tl * gridDim.x) * g.a_bytes0;
        for (int q = 0; q < 4; q++) {
          const int s0 = (q * ks) / 4, s1 = ((q + 1) * ks) / 4;
          if (s1 > s0) {
            tc::mbar_expect_tx(&a_part[buf][q], (uint32_t)(s1 - s0) * 4096);
            tc::bulk_g2s(sa + (size_t)s0 * 4096, src + (size_t)s0 * 4096, (uint32_t)(s1 - s0) * 4096, &a_part[buf][q]);
          } else {
            tc::mbar_arrive(&a_part[buf][q]);
          }
        }
      }
    }
  } else if (warp == 5) {
    // ---- weight loader: the slabs of every layer in MMA order, `ring` slots ahead (converged warp, one elected lane issues) ------
    const bool leader = tc::elect_one();
    int r = 0, gg = 0;
    uint32_t rpar = 1;
    for (int tl = 0; tl < T; tl++)
      for (int l = 0; l < g.nl; l++) {
        const int slab = g.nb[l] * 32, ks = g.ks[l];
        for (int nb = 0; nb < g.nblk[l]; nb++)
          for (int s0 = 0; s0 < ks; s0 += CH_GG, gg++) {
            const int cnt = min(CH_GG, ks - s0);
            if (gg >= g.ring && !failed && !tc::mbar_wait<cTF_L>(&ring_empty[r], rpar)) failed = true;
            if (s0 == 0) CTR(2, g.blk0[l] + nb, 0);
            if (leader) {
              tc::mbar_expect_tx(&ring_full[r], (uint32_t)(cnt * slab));
              tc::bulk_g2s(ring + (size_t)r * CH_SLOT_BYTES, g.w4s[l] + (size_t)(nb * ks + s0) * slab, (uint32_t)(cnt * slab), &ring_full[r]);
            }
            if (++r == g.ring) { r = 0; rpar ^= 1u; }
          }
      }
  } else if (warp == 6) {
    // ---- issuer ---------------------------------------------------------------------------------------------------------------
    const bool leader = tc::elect_one();
    const uint32_t d_hi = tc::desc_hi(128);
    const uint32_t idesc128 = tc::make_idesc_mxf4(128, 128), idesc64 = tc::make_idesc_mxf4(128, 64);
    int r = 0;
    uint32_t rpar = 0;
    for (int tl = 0; tl < T; tl++) {
      const int buf = tl % g.nbuf0, use = tl / g.nbuf0;
      for (int l = 0; l < g.nl; l++) {
        const int NBl = g.nb[l], ks = g.ks[l], slab16 = NBl * 2;
        const uint32_t a_lo = tc::desc_lo(sa_u0 + (uint32_t)g.a_off[l] + (l == 0 ? (uint32_t)buf * g.a_bytes0 : 0u), 2048);
        const uint32_t b_lo0 = tc::desc_lo(ring_u, (uint32_t)NBl * 16);
        const uint32_t idesc = NBl == 128 ? idesc128 : idesc64;
        const uint32_t accl = is_acc(l) ? 1u : 0u;
        int qn = 0;
        for (int nb = 0; nb < g.nblk[l]; nb++) {
          const int gb = tl * nbt + g.blk0[l] + nb, slot = gb % SLOTS, u = gb / SLOTS;
          CTR(0, g.blk0[l] + nb, 0);
          if (u >= 1 && !failed && !tc::mbar_wait<cTF_I>(&empty[slot], (uint32_t)(u - 1) & 1)) failed = true;
          tc::fence_after_sync();
          CTR(0, g.blk0[l] + nb, 1);
          const uint32_t acc = tmem + (uint32_t)slot * 128;
          for (int s0 = 0; s0 < ks; s0 += CH_GG) {
            if (!failed && !tc::mbar_wait<cTF_I>(&ring_full[r], rpar)) failed = true;
            if (s0 == 0) CTR(0, g.blk0[l] + nb, 2);
            const int cnt = min(CH_GG, ks - s0);
            const uint32_t b_lo = b_lo0 + (uint32_t)r * (CH_SLOT_BYTES / 16);
            const uint32_t a_s0 = a_lo + (uint32_t)(2 * s0) * 128;
            auto block = [&](int n, int e) {  // n MMAs in one asm block, from slab s0 + e of this group
              const uint32_t aa = a_s0 + (uint32_t)(2 * e) * 128, bb = b_lo + (uint32_t)e * slab16, first = (accl || s0 + e > 0) ? 1u : 0u;
              if (!leader) return;
              if (NBl == 128) {
                if (n == 8) tc::mma8_mxf4<256, 256>(acc, aa, bb, d_hi, idesc, sf, first);
                else if (n == 4) tc::mma4_mxf4<256, 256>(acc, aa, bb, d_hi, idesc, sf, first);
                else tc::mma2_mxf4<256, 256>(acc, aa, bb, d_hi, idesc, sf, first);
              } else {
                if (n == 8) tc::mma8_mxf4<256, 128>(acc, aa, bb, d_hi, idesc, sf, first);
                else if (n == 4) tc::mma4_mxf4<256, 128>(acc, aa, bb, d_hi, idesc, sf, first);
                else tc::mma2_mxf4<256, 128>(acc, aa, bb, d_hi, idesc, sf, first);
              }
            };
            if (nb > 0 && (cnt == 8 || cnt == 4)) {  // every operand slab of the tile is in: one asm block for the group
              block(cnt, 0);
            } else if (nb == 0 && l > 0 && (cnt & 1) == 0) {
              // the tile's first block runs behind the layer below: slabs 2b, 2b+1 = the 128 channels of its block b
              for (int e = 0; e < cnt; e += 2) {
                if (!failed && !tc::mbar_wait<cTF_I>(&a_ready[l][(s0 + e) >> 1], (uint32_t)tl & 1)) failed = true;
                tc::fence_after_sync();
                block(2, e);
              }
            } else if (nb == 0 && l == 0 && (cnt == 8 || cnt == 4)) {
              // the chain's first operand tile arrives in K quarters: the quarters this group reads, then one asm block
              while (qn < 4 && s0 + cnt - 1 >= (qn * ks) / 4) {
                if (!failed && !tc::mbar_wait<cTF_I>(&a_part[buf][qn], (uint32_t)use & 1)) failed = true;
                qn++;
              }
              tc::fence_after_sync();
              block(cnt, 0);
            } else {
              for (int s = s0; s < s0 + cnt; s++) {
                if (nb == 0) {  // the first operand tile arrives in K quarters
                  if (l == 0) {
                    while (qn < 4 && s >= (qn * ks) / 4) {
                      if (!failed && !tc::mbar_wait<cTF_I>(&a_part[buf][qn], (uint32_t)use & 1)) failed = true;
                      qn++;
                    }
                  } else if ((s & 1) == 0) {
                    if (!failed && !tc::mbar_wait<cTF_I>(&a_ready[l][s >> 1], (uint32_t)tl & 1)) failed = true;
                  }
                  tc::fence_after_sync();
                }
                if (leader)
                  tc::mma_mxf4_lh(acc, a_lo + (uint32_t)(2 * s) * 128, d_hi, b_lo + (uint32_t)(s - s0) * slab16, d_hi, idesc, sf, sf, (accl || s > 0) ? 1u : 0u);
              }
            }
            if (leader) tc::commit(&ring_empty[r]);
            if (++r == g.ring) { r = 0; rpar ^= 1u; }
            __syncwarp();
          }
          if (leader) tc::commit(&full[slot]);
          __syncwarp();
          CTR(0, g.blk0[l] + nb, 3);
        }
        if (l == g.rel_layer) {
          if (leader) tc::commit(&a_empty[buf]);
          __syncwarp();
        }
      }
    }
  } else if (warp < 4) {
    // ---- epilogue: one image (row) per thread -----------------------------------------------------------------------------------
    const int row = hw * 32 + lane;
    for (int tl = 0; tl < T; tl++) {
      const long long img = ((long long)blockIdx.x + (long long)tl * gridDim.x) * 128 + row;
      int cls_max = 0, cls_arg = 0;
      for (int l = 0; l < g.nl; l++) {
        const float* st = sthr + g.thr_off[l];
        for (int nb = 0; nb < g.nblk[l]; nb++) {
          const int b = g.blk0[l] + nb, gb = tl * nbt + b, slot = gb % SLOTS, u = gb / SLOTS;
          if (warp == 0) CTR(1, b, 0);
          if (!failed && !tc::mbar_wait<cTF_E>(&full[slot], (uint32_t)u & 1)) failed = true;
          tc::fence_after_sync();
          if (warp == 0) CTR(1, b, 1);
          const uint32_t trow = tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * 128;
          uint4 c4[4];
          uint32_t word[2] = {0, 0};
          if (l < last) {
            // hidden layer: 128 channels -> four 16-byte rows of e2m1 codes
            uint32_t hh[2][16];
            tc::tmem_ld32_pack16(trow, hh[0]);
#pragma unroll
            for (int j = 0; j < 4; j++) {
              uint32_t dl[16];
              if (TERN) {
                const uint4* dp = reinterpret_cast<const uint4*>(reinterpret_cast<const uint32_t*>(st) + g.N[l] + (nb * 128 + j * 32) / 2);
#pragma unroll
                for (int q = 0; q < 4; q++) { const uint4 v = dp[q]; dl[4 * q] = v.x; dl[4 * q + 1] = v.y; dl[4 * q + 2] = v.z; dl[4 * q + 3] = v.w; }
              }
              tc::tmem_ld_wait_dep16(hh[j & 1]);
              if (j + 1 < 4) tc::tmem_ld32_pack16(trow + (j + 1) * 32, hh[(j + 1) & 1]);
              if (TERN) c4[j] = tern_codes16(hh[j & 1], dl);
              else c4[j] = bin_codes16(hh[j & 1]);
            }
          } else if (LASTM == OUT_SCORE) {
            // PassThroughActivation<ap_uint<16>>: (S + offset) mod 2^16; 64 scores per image, class = first score strictly above
            // a running max from 0 over the first ncls (foldedmv-offload.h:398-406)
            uint32_t r[2][32];
            tc::tmem_ld32(trow, r[0]);
#pragma unroll
            for (int j = 0; j < 2; j++) {
              const float* th = st + j * 32;
              tc::tmem_ld_wait_dep(r[j & 1]);
              if (j + 1 < 2) tc::tmem_ld32(trow + (j + 1) * 32, r[(j + 1) & 1]);
              if (img < n_img) {
                int16_t* o16 = reinterpret_cast<int16_t*>(out) + (size_t)img * 64 + j * 32;
#pragma unroll
                for (int c = 0; c < 32; c += 8) {
                  uint32_t pk[4];
#pragma unroll
                  for (int e = 0; e < 4; e++) {
                    const uint32_t lo = (uint32_t)(__float2int_rn(__uint_as_float(r[j & 1][c + 2 * e])) + __float2int_rn(th[c + 2 * e])) & 0xffff;
                    const uint32_t hi = (uint32_t)(__float2int_rn(__uint_as_float(r[j & 1][c + 2 * e + 1])) + __float2int_rn(th[c + 2 * e + 1])) & 0xffff;
                    pk[e] = lo | (hi << 16);
                    const int ch = j * 32 + c + 2 * e, vlo = (int)(int16_t)lo, vhi = (int)(int16_t)hi;
                    if (ch < g.ncls && vlo > cls_max) { cls_max = vlo; cls_arg = ch; }
                    if (ch + 1 < g.ncls && vhi > cls_max) { cls_max = vhi; cls_arg = ch + 1; }
                  }
                  *reinterpret_cast<uint4*>(o16 + c) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                }
              }
            }
            if (g.classes && img < n_img) g.classes[img] = cls_arg;
          } else {
            // LFC: the 64 decision bits of the output layer
            uint32_t hh[2][16];
            tc::tmem_ld32_pack16(trow, hh[0]);
            tc::tmem_ld32_pack16(trow + 32, hh[1]);
            tc::tmem_ld_wait_dep16(hh[0]);
            tc::tmem_ld_wait_dep16(hh[1]);
            word[0] = tern_bit_word(hh[0]);
            word[1] = tern_bit_word(hh[1]);
          }
          if (warp == 0) CTR(1, b, 2);
          if (gb + SLOTS < T * nbt) {  // this slot's next user
            const int b2 = (b + SLOTS) % nbt;
            if (is_acc(blk_layer[b2])) preload(slot, blk_layer[b2], blk_nb[b2]);
          }
          tc::fence_before_sync();
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(&empty[slot]);  // this warp's quarter of the accumulator has been read (and re-loaded)
          if (warp == 0) CTR(1, b, 3);
          if (l < last) {
            uint4* dst = reinterpret_cast<uint4*>(smem + g.a_off[l + 1]) + (size_t)(nb * 4) * 128 + row;
#pragma unroll
            for (int j = 0; j < 4; j++) dst[(size_t)j * 128] = c4[j];
            tc::fence_async_smem();  // generic-proxy stores, read by the tensor core's async proxy
            tc::mbar_arrive(&a_ready[l + 1][nb]);
            if (warp == 0) CTR(1, b, 4);
          } else if (LASTM != OUT_SCORE && img < n_img) {
            // the class of the 64-bit word, multi-image rule (foldedmv-offload.cpp:202-220; same arithmetic as k_lfc_classes)
            if (g.classes) {
              const uint64_t w = (((uint64_t)word[1] << 32) | word[0]) & (g.ncls >= 64 ? ~0ull : (g.ncls <= 0 ? 0ull : (~0ull >> (64 - g.ncls))));
              g.classes[img] = !w ? 0 : (w < (1ull << 32) ? 31 - __clz((unsigned)w) : (int)(unsigned)log2((double)w));
            }
            *reinterpret_cast<uint2*>(out + (size_t)img * 2) = make_uint2(word[0], word[1]);
          }
        }
      }
    }
  }
  if (failed && status) *status = 1;
  tc::fence_before_sync();
  __syncthreads();
#ifdef BNNB_CHAIN_TRACE
  if (blockIdx.x == 0 && tid == 0 && T > TRT) {
    const long long t0 = tr[0][0][0];
    for (int b = 0; b < nbt && b < 16; b++)
      printf("CHAIN b%d L%d: issuer slot-wait %lld slot %lld ring %lld committed %lld | epilogue wait %lld full %lld codes %lld released %lld handed %lld | loader %lld\n", b,
             (int)blk_layer[b], tr[0][b][0] - t0, tr[0][b][1] - t0, tr[0][b][2] - t0, tr[0][b][3] - t0, tr[1][b][0] - t0, tr[1][b][1] - t0, tr[1][b][2] - t0,
             tr[1][b][3] - t0, tr[1][b][4] - t0, tr[2][b][0] - t0);
  }
#endif
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

}  // namespace

// Shared-memory plan: the first operand tile (twice when it fits, so that the next tile's copy runs under this tile's layers), the
// tiles of odd and of even layers in two regions of their own - a layer's epilogue writes the next tile while its own is still
// being read - then the weight ring with whatever is left (at least two slots) and the threshold tables.  When that does not fit,
// the even layers' region aliases the first operand buffer (single), which is then released after the last even layer.
static bool plan_chain(ChainGeom& g, int* smem_bytes) {
  const int budget = 226 * 1024;
  int a[CH_MAX], thr = 0;
  for (int l = 0; l < g.nl; l++) {
    a[l] = g.K[l] / 32 * 2048;
    g.thr_off[l] = thr / 4;
    thr += (g.N[l] + g.N[l] / 2 + 127) / 128 * 128 * 4;
  }
  int P = 0, Q = 0, last_even = 0;
  for (int l = 1; l < g.nl; l++) {
    if (l & 1) P = std::max(P, a[l]);
    else { Q = std::max(Q, a[l]); last_even = l; }
  }
  g.a_bytes0 = a[0];
  int best = -1, best_ring = 0;
  for (int opt = 0; opt < 3; opt++) {
    const int nbuf0 = opt == 0 ? 2 : 1;
    const bool alias = opt == 2;
    if (alias && (Q > a[0] || Q == 0)) continue;
    const int fixed = nbuf0 * a[0] + P + (alias ? 0 : Q) + thr;
    const int ring = (budget - fixed) / CH_SLOT_BYTES;
    if (ring >= 3) { best = opt; best_ring = ring; break; }
    if (ring > best_ring) { best = opt; best_ring = ring; }
  }
  if (best < 0 || best_ring < 2) return false;
  g.nbuf0 = best == 0 ? 2 : 1;
  const bool alias = best == 2;
  g.rel_layer = alias ? last_even : 0;
  g.a_off[0] = 0;
  const int p_off = g.nbuf0 * a[0], q_off = alias ? 0 : p_off + P;
  for (int l = 1; l < g.nl; l++) g.a_off[l] = (l & 1) ? p_off : q_off;
  g.off_ring = p_off + P + (alias ? 0 : Q);
  g.ring = std::min(best_ring, CH_RING_MAX);
  g.off_thr = g.off_ring + g.ring * CH_SLOT_BYTES;
  *smem_bytes = g.off_thr + thr;
  return true;
}

// Layers L[0..nl) in one launch.  Returns false when the run does not qualify (the caller launches the layers one by one):
// every layer must be a gemm_f4 layer, the first one fed with ACT_F4T tiles, hidden layers of one output kind (1-bit or 2-bit) with
// a multiple of 128 channels, the last one the network's output layer (64 channels: 16-bit scores or the LFC output word).
bool launch_fc_chain(const DevLayer* L, int nl, const uint32_t* in, uint32_t* out, int n_img, int num_sms, int in_fmt, int* status, cudaStream_t s,
                     int* classes, int ncls) {
  if (nl < 2 || nl > CH_MAX || n_img <= 0 || in_fmt != ACT_F4T) return false;
  ChainGeom g{};
  g.nl = nl;
  const int hidden = L[0].d.out_mode;
  if (hidden != OUT_BIT && hidden != OUT_TERN) return false;
  int blk = 0;
  for (int l = 0; l < nl; l++) {
    const LayerDesc& d = L[l].d;
    if (!gemm_f4_supports(L[l])) return false;
    if (d.in_mode != (l == 0 ? d.in_mode : (hidden == OUT_TERN ? IN_TERN : IN_XNOR))) return false;  // the codes the layer below writes
    if (l == 0 && d.in_mode != IN_XNOR && d.in_mode != IN_TERN) return false;
    if (l > 0 && d.K != L[l - 1].d.N) return false;
    if (l < nl - 1) {
      if (d.out_mode != hidden || d.N % 128 || d.N / 128 > 8) return false;
    } else if (d.N != 64 || (d.out_mode != OUT_SCORE && d.out_mode != OUT_BIT)) return false;
    g.K[l] = d.K; g.N[l] = d.N; g.ks[l] = d.K / 64;
    g.nb[l] = l < nl - 1 ? 128 : 64;
    if (g.nb[l] != gemm_f4_block(d)) return false;  // the slabs k_repack_w4s laid out
    g.nblk[l] = d.N / g.nb[l];
    g.blk0[l] = blk;
    blk += g.nblk[l];
    g.w4s[l] = L[l].w4s;
    g.thr_tc[l] = L[l].thr_tc;
  }
  if (blk > CH_BLK_MAX) return false;
  g.nblk_tile = blk;
  g.classes = classes;
  g.ncls = ncls < 0 ? 0 : (ncls > 64 ? 64 : ncls);
  int smem = 0;
  if (!plan_chain(g, &smem)) return false;
  const int ntiles = (n_img + 127) / 128;
  const int grid = ntiles < num_sms ? ntiles : num_sms;
  const int lastm = L[nl - 1].d.out_mode;
  auto go = [&](auto kern) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    launch_maybe_pdl(kern, grid, 256, (size_t)smem, s, g, reinterpret_cast<const uint8_t*>(in), out, n_img, status);
  };
  if (hidden == OUT_TERN) { if (lastm == OUT_SCORE) go(k_fc_chain<true, OUT_SCORE>); else go(k_fc_chain<true, OUT_BIT>); }
  else { if (lastm == OUT_SCORE) go(k_fc_chain<false, OUT_SCORE>); else go(k_fc_chain<false, OUT_BIT>); }
  return true;
}

}  // namespace bnnb
