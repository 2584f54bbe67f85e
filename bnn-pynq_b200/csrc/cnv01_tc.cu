// Fused CNV layers 0+1: int8 image -> conv3x3 (3->64, tcgen05.mma kind::i8) -> threshold -> conv3x3 (64->64, kind::mxf4 with unit
// scales, or kind::i8 on core 1) -> threshold -> 2x2 max-pool -> 14x14x64 output as e2m1 operand planes (ACT_F4P) or bit-packed.
// Layer 0's activations never exist in bit-packed form or in global memory: its epilogue writes them as e2m1 nibbles (int8 on core 1)
// straight into layer 1's UMMA operand planes in shared memory.  Both layers' thresholds ride in the accumulators (constant operand
// plane for layer 0, biased pre-load for layer 1) and both epilogues read 16 bits per channel (see FOLD / TFOLD in the kernel).
//
// Replaces (cnvW1A1/hw/top.cpp:210-217): Mem2Stream + DataWidthConverters + ConvLayer_Batch L0 + ConvLayer_Batch L1 +
// StreamingMaxPool_Batch.
//
// Layer 0 as an implicit GEMM (K = 27): one operand plane P with 16 bytes per row, row m = bytes [3m, 3m+16) of the HWC
// image (m = y*32 + x: the 3x3x3 window row ky of pixel (y,x) is the first 9 bytes of P[m + 32*ky]).  K is laid out as
// [ky][16]: k = ky*16 + (kx*3 + c) with zero weights for the 7 trailing bytes of each ky and for the whole 4th chunk, so
// one tile needs two K=32 MMAs whose A descriptors differ only by a start shift (LBO = 32 rows).
// Warp-specialised CTA (one per SM, 736 threads = 23 warps): eight warps run layer 1's epilogue on image i, eight warps run layer 0's
// epilogue of image i+1 into the other layer-1 operand buffer, three issuer warps (two for layer 1, one for layer 0 + the TMA bulk
// prefetch of images two ahead), four helper warps (window planes, max-pool + store).  See the role table above the kernel.
// TAP (template flag, parity tests only): the same kernel additionally stores every layer-0 / layer-1 MMA sum and layer 0's output
// operand rows to global memory (bnnb_tap on the tensor cores); the production instantiation carries none of that code.
#include <climits>
#include <cstdio>
#include <type_traits>
#include "kernels.cuh"
#include "tcgen05.cuh"
#include "f4_common.cuh"

namespace bnnb {
namespace {

constexpr int L1_PITCH = 960;                 // 7 tiles * 128 + max shift 62, rounded to 8
constexpr int L1_PLANE = L1_PITCH * 16;       // bytes per 16-channel plane
// layer 1 runs either on kind::i8 (16 channels per 16-byte plane row, 4 planes, 18 K=32 MMAs per tile) or on kind::mxf4 with
// unit scales (F4: 32 e2m1 channels per plane row, 2 planes, 9 K=64 MMAs per tile: half the operand bytes, twice the MAC rate)
constexpr int P_ROWS = 1120;                  // 8 tiles * 128 + 64 (ky = 2) + 32 (dummy chunk)
constexpr int P_BYTES = P_ROWS * 16;
constexpr int IMG_BYTES = 3072 + 64;          // image + slack for the 16-byte windows of the last rows
constexpr int B0_BYTES = 4 * 64 * 16;         // layer-0 weights: 4 chunks x 64 rows x 16 bytes
constexpr int L1_TILES = 7, L0_TILES = 8;
#ifndef BNNB_PLANEU
#define BNNB_PLANEU 1
#endif
#ifndef BNNB_POOLU
#define BNNB_POOLU 1
#endif
#ifndef BNNB_NH
#define BNNB_NH 4
#endif
#ifdef BNNB_NOEPI
constexpr bool kEpi = false;  // experiment: epilogue bodies compiled out (results wrong, hand-offs kept)
#else
constexpr bool kEpi = true;
#endif
#ifdef BNNB_NOMMA
constexpr bool kMma = false;  // experiment: MMAs compiled out (commits kept)
#else
constexpr bool kMma = true;
#endif
#ifdef BNNB_NOMMA0
constexpr bool kMma0 = false;  // experiment: layer-0 MMAs only compiled out
#else
constexpr bool kMma0 = kMma;
#endif
#ifdef BNNB_NOMMA1
constexpr bool kMma1 = false;  // experiment: layer-1 MMAs only compiled out
#else
constexpr bool kMma1 = kMma;
#endif
#ifdef BNNB_NOPLANE
constexpr bool kPlane = false;  // experiment: window-plane build compiled out (hand-offs kept)
#else
constexpr bool kPlane = true;
#endif
#ifdef BNNB_NOPOOL
constexpr bool kPool = false;   // experiment: max-pool + store compiled out (hand-offs kept)
#else
constexpr bool kPool = true;
#endif
constexpr int PLANEU = BNNB_PLANEU;           // unroll of the window-plane loop
constexpr int POOLU = BNNB_POOLU;             // unroll of the pool loop
constexpr int NH = BNNB_NH;                   // helper warps (window planes, max-pool + store)
constexpr int NT01 = (8 + 8 + 3 + NH) * 32;   // threads per CTA

template <int OUTM, bool F4> struct Smem01 {
  static constexpr int OPL = OUTM == OUT_TERN ? 2 : 1;
  static constexpr int NPL = F4 ? 2 : 4;               // operand planes per buffer
  static constexpr int L1_A = NPL * L1_PLANE;          // one operand buffer (64 channels)
  static constexpr int L1_B = 9 * NPL * 64 * 16;       // weights 64 x 576
  static constexpr int BITS = L1_TILES * 128 * 2 * OPL * 4;
  static constexpr int NBITS = F4 ? 2 : 1;             // bit-row buffers (the int8 variant has no room for a second one)
  // FOLD variant (fp4, 1-bit outputs): each window plane is followed by a plane of constant rows {64, 1, 0, ...}: the fourth K chunk of
  // layer 0, against which the weights carry the thresholds (thr = 64*a + b), so that the MMAs leave thr - S without any pre-load
  static constexpr bool FOLD = F4 && OUTM == OUT_BIT;
  static constexpr bool TFOLD = F4 && OUTM == OUT_TERN;  // the 2-bit counterpart (see the kernel)
  static constexpr int P_STRIDE = F4 ? 2 * P_BYTES : P_BYTES;
  static constexpr int OFF_A = 0, OFF_B1 = 2 * L1_A, OFF_P = OFF_B1 + L1_B, OFF_IMG = OFF_P + 2 * P_STRIDE, OFF_B0 = OFF_IMG + 2 * IMG_BYTES,
                       OFF_BITS = OFF_B0 + B0_BYTES, OFF_THR = OFF_BITS + NBITS * BITS, TOTAL = OFF_THR + 2 * 64 * 2 * 4;
};

// FOLD epilogue of layer 0 (below): operand nibble 2k of every 8-channel word holds channel k, nibble 2k+1 channel 4+k; layer 1's
// weights are put in the same order
__device__ __forceinline__ uint32_t fold_nibble_order(uint32_t w) {
  uint32_t o = 0;
#pragma unroll
  for (int n = 0; n < 8; n++) o |= ((w >> (4 * ((n >> 1) + 4 * (n & 1)))) & 0xFu) << (4 * n);
  return o;
}
// bytes 1 and 3 of a and of b, each replaced by its sign bit replicated (0x00 / 0xFF): four 16-bit signs -> four byte masks
__device__ __forceinline__ uint32_t sign_bytes(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, 0xFDB9;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
// TFOLD: accumulator column p of layer 1 holds channel (p & 32) + 8 * (p & 3) + ((p & 31) >> 2): columns 4g .. 4g+3 = channels g, 8+g,
// 16+g, 24+g of the 32-channel half, so that one PRMT over two .pack::16b registers yields the four bytes of a 32-channel bit word
__device__ __forceinline__ int tfold_channel(int p) { return (p & 32) + 8 * (p & 3) + ((p & 31) >> 2); }
__device__ __forceinline__ void half_sync(int id) { asm volatile("bar.sync %0, 256;" ::"r"(id) : "memory"); }

// Warp roles (736 threads): warps 0-7 = half A (layer-1 epilogue, pool, store), warps 8-15 = half B (layer-0 epilogue into the
// other operand buffer), each half split into two teams of four warps by 32-channel half (the per-tile epilogue is a latency
// chain: two teams halve it); warps 16-17 = layer-1 MMA issuers, warp 18 = image prefetch (TMA) + layer-0 MMA issuer, warps 19-22 = helpers
// (window planes two images ahead, 2x2 max-pool of the bit rows + store).
// The roles are decoupled by mbarriers only (no CTA-wide barrier in steady state):
//   img_full[b]   TMA    -> half A   image buffer b holds an image                      (expect_tx 3072 bytes)
//   p_full[b]     half A -> warp 16   window plane b is built (and image buffer b free)  (256 arrivals)
//   p_empty[b]    warp 16 -> half A   all layer-0 MMAs reading plane b have completed    (tcgen05.commit)
//   f0[s], e0[s]  warp 16 <-> half B  layer-0 accumulator slot s (of 2) full / read out   (commit / 8 arrivals)
//   l0_done[b][t] half B -> warps 16-17  layer-0 tile t of the image in operand buffer b is written (256 arrivals): layer-1 tile t-1 may start
//   abuf_empty[b] warp 16 -> half B   all layer-1 MMAs reading buffer b have completed   (tcgen05.commit)
//   full[s], empty[s] warp 16 <-> half A  layer-1 accumulator slot s (of 4) full / read out
template <int OUTM, bool F4, bool TAP>  // OUT_BIT: 1-bit activations (W1A1); OUT_TERN: ternary activations (W1A2 / W2A2)
__global__ void __launch_bounds__(NT01, 1) k_cnv01_tc(DevLayer L0, DevLayer L1, const int8_t* __restrict__ in, uint32_t* __restrict__ out,
                                                      int n_img, int out_fmt, int* __restrict__ status, Tap01 tp) {
  tc::griddep_launch();  // dependents (programmatic dependent launch) may start their prologues now
  using S = Smem01<OUTM, F4>;
  // test-first mbarrier waits (tcgen05.cuh), by site: bit 0 img_full, 1 p_empty, 2 bits_full (helper warps); 3 abuf_empty, 4 f0 (layer-0
  // epilogue); 5 p_full, 6 e0, 7 empty, 8 l0_done (issuers); 9 bits_empty, 10 full (layer-1 epilogue)
#ifndef BNNB_01_TFS
#define BNNB_01_TFS 0x1f8
#endif
#ifndef BNNB_01_TFS_T
#define BNNB_01_TFS_T 0
#endif
  constexpr int TFSM = OUTM == OUT_BIT ? BNNB_01_TFS : BNNB_01_TFS_T;
#define TFS(n) (((TFSM >> (n)) & 1) != 0)
  constexpr int L1_A = S::L1_A;
  constexpr uint32_t SF_COL = 448;  // TMEM columns of the unit scale factors (F4)
  constexpr uint32_t L0_ACC = 256;  // TMEM columns: layer-1 slots 0,64,128,192; layer-0 slots 256,320
#ifndef BNNB_L0S
#define BNNB_L0S 2
#endif
  // layer-0 accumulator slots (TMEM columns 256, 320, ...).  A tile's round trip (issue -> MMAs -> hand-off -> epilogue -> hand-off)
  // is ~1000 clocks.  Round-2 measurements: the kernel is bound by this hand-off skeleton, not by the tensor pipe (all
  // MMAs compiled out: 1.19 of 1.35 ms); more slots (3, 4, 5), teams on alternate tiles, one polling warp per group and one arrival
  // per warp were all SLOWER - whatever lengthens the path layer-0 accumulator -> l0_done costs more than it frees.
  constexpr int L0_SLOTS = BNNB_L0S;
  // FOLD: the thresholds are folded into the accumulators, arranged so that the SIGN BIT of the result is the decision thr < acc.
  // Layer 0: the thresholds ride in the MMA itself - the fourth K chunk of the second MMA reads a plane of constant rows {64, 1, 0..}
  // against weight columns (a, b) with 64*a + b = thr[n] (clamped to the reachable range, so the result fits 16 bits) - and the
  // MMAs run on negated weights: the accumulator is thr - S, negative <=> decision 1, with no pre-load to write and wait for in the
  // epilogue chain (round 2: 1.39 -> 1.35 ms).  The epilogue reads the accumulators with .pack::16b (two channels per
  // register), turns four signs into four 0x00 / 0xFF bytes with one PRMT and merges two such words into eight operand nibbles
  // with one LOP3: three ALU instructions per eight channels.  A set decision is therefore the e2m1 code 0xF = -6.0, not 1.0:
  // layer 1 accumulates -6 * (its true sum) onto a pre-load of 6 * thr + 3 (exact in fp32), negative <=> thr < sum, and its
  // epilogue is one funnel shift per channel instead of compare + select + insert.
  constexpr bool FOLD = F4 && OUTM == OUT_BIT;
  // TFOLD (fp4, 2-bit outputs): the same machinery carries the FIRST threshold, biased so that the low 16 bits of every accumulator
  // are U = thr0 - S + 32768 in [0, 65536) - layer 0 through the constant plane {64, 1, 127, 127} (64a + b + 2 * 127 * 127 = thr0 +
  // 32768), layer 1 through a pre-load of 1.5 * 2^23 + 32768 + thr0 (the float's low mantissa bits ARE the integer; the tensor core's
  // fp32 accumulation is exact down to the last mantissa bit, measured) and negated weights.  The epilogues read two channels per
  // register (.pack::16b); bit 15 of a half is NOT (thr0 < S), and ONE packed add of (thr1 - thr0) per channel pair gives the same
  // for the second threshold (no half leaves [0, 65536), so nothing carries).  PRMT turns four such bits into byte masks: twelve
  // ALU instructions per eight channels instead of thirty-two, and no threshold is fetched from shared memory per tile.
  constexpr bool TFOLD = F4 && OUTM == OUT_TERN;
  constexpr bool CP = FOLD || TFOLD;  // thresholds ride in the accumulators: constant plane, negated sums, permuted operand nibbles
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[4], empty[4], l0_done[2][L0_TILES], abuf_empty[2], f0[L0_SLOTS], e0[L0_SLOTS], p_full[2], p_empty[2], img_full[2], bits_full[2], bits_empty[2],
      pre_done;
  __shared__ uint32_t tmem_base;
  uint8_t* sB1 = smem + S::OFF_B1;
  uint8_t* sP = smem + S::OFF_P;
  uint8_t* sImg = smem + S::OFF_IMG;
  uint8_t* sB0 = smem + S::OFF_B0;
  uint32_t* sbits = reinterpret_cast<uint32_t*>(smem + S::OFF_BITS);
  int32_t* sthr0 = reinterpret_cast<int32_t*>(smem + S::OFF_THR);  // layer 0: floor(thr/2)   [64][2]
  int32_t* sthr1 = sthr0 + 128;                                    // layer 1: adjusted       [64][2] (F4: as float)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int role = warp < 8 ? 0 : (warp < 16 ? 1 : (warp < 19 ? 2 : 3));             // 0: half A, 1: half B, 2: issuers (warps 16, 17: layer 1; warp 18: layer 0 + TMA)
  const int ht = tid & 255, hw = warp & 3, team = (warp >> 2) & 1;  // thread in the half; TMEM lane quarter; 32-channel half

  // ---- one-time setup -------------------------------------------------------------------------------------------------
  for (int i = tid; i < 64 * 9 * S::NPL; i += NT01) {  // layer-1 weights -> chunk-major operand layout
    const int p = i % 64, kc = i / 64, n = CP ? tfold_channel(p) : p;  // accumulator column p holds channel n
    const uint8_t* src = F4 ? L1.w4 + (size_t)n * (L1.kpad / 2) : reinterpret_cast<const uint8_t*>(L1.w8) + (size_t)n * L1.kpad;
    uint4 wv = *reinterpret_cast<const uint4*>(src + kc * 16);
    if (CP) { wv.x = fold_nibble_order(wv.x); wv.y = fold_nibble_order(wv.y); wv.z = fold_nibble_order(wv.z); wv.w = fold_nibble_order(wv.w); }
    if (TFOLD) { wv.x ^= 0x88888888u; wv.y ^= 0x88888888u; wv.z ^= 0x88888888u; wv.w ^= 0x88888888u; }  // e2m1 sign bits: the MMAs subtract
    *reinterpret_cast<uint4*>(sB1 + ((size_t)kc * 64 + p) * 16) = wv;
  }
  for (int i = tid; i < 64 * 64; i += NT01) {  // layer-0 weights: k = ky*16 + j, j = kx*3 + c < 9; L0.wp holds int8x4 words [64][8], k natural
    const int n = i / 64, k = i % 64, ky = k / 16, j = k % 16;
    int8_t v = 0;
    if (ky < 3 && j < 9) v = reinterpret_cast<const int8_t*>(L0.wp)[n * 32 + ky * 9 + j];
    if (CP && ky == 3 && j < (TFOLD ? 4 : 2)) {  // threshold columns against the constant plane {64, 1, 127, 127}
      const int t = L0.thr[2 * n];   // |S| <= 27 * 128: thresholds outside [-3457, 3456] decide like the nearest bound
      int th = t == INT_MAX ? 3456 : max(min(t >> 1, 3456), -3457);  // thr < 2S  <=>  floor(thr/2) < S
      if (TFOLD) th += 32768 - 2 * 127 * 127;                        // 64*a + b + 127*127 + 127*127 = thr + 32768
      sB0[(size_t)3 * (64 * 16) + n * 16 + j] = (uint8_t)(int8_t)(j == 0 ? (th >> 6) : (j == 1 ? (th & 63) : 127));
      continue;
    }
    sB0[(size_t)(k / 16) * (64 * 16) + n * 16 + (k % 16)] = (uint8_t)(CP ? -v : v);  // the MMAs subtract (see above)
  }
  for (int i = tid; i < 128; i += NT01) {
    const int t = L0.thr[i];
    sthr0[i] = t == INT_MAX ? (1 << 26) : min(t >> 1, 1 << 26);  // thr < 2S  <=>  floor(thr/2) < S; |S| <= 3456, and thr - S must not overflow
    const int t1 = L1.thr_tc[i];
    if (F4) reinterpret_cast<float*>(sthr1)[i] = t1 == INT_MAX ? 3.0e38f : (float)t1;
    else sthr1[i] = t1;
  }
  for (int i = tid; i < 2 * L1_A / 16; i += NT01) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);  // incl. pad pixels
  for (int i = tid; i < (2 * S::P_STRIDE + 2 * IMG_BYTES) / 16; i += NT01) reinterpret_cast<uint4*>(sP)[i] = make_uint4(0, 0, 0, 0);
  if (CP) {  // the constant planes
    __syncthreads();
    for (int i = tid; i < 2 * P_ROWS; i += NT01)
      reinterpret_cast<uint4*>(sP + (size_t)(i / P_ROWS) * S::P_STRIDE + P_BYTES)[i % P_ROWS] = make_uint4(0x7F7F0140u, 0, 0, 0);  // bytes 64, 1, 127, 127, 0, ...
  }
  if (warp == 0) tc::tmem_alloc(&tmem_base, 512);
  if (tid == 0) {
    for (int i = 0; i < 2; i++) {
      tc::mbar_init(&abuf_empty[i], 2);
      for (int t = 0; t < L0_TILES; t++) tc::mbar_init(&l0_done[i][t], 256);
      tc::mbar_init(&p_full[i], NH * 32); tc::mbar_init(&bits_full[i], 8); tc::mbar_init(&bits_empty[i], NH); tc::mbar_init(&p_empty[i], 1); tc::mbar_init(&img_full[i], 1);
    }
    for (int i = 0; i < 4; i++) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 8); }
    for (int i = 0; i < L0_SLOTS; i++) { tc::mbar_init(&f0[i], 1); tc::mbar_init(&e0[i], 8); }
    tc::mbar_init(&pre_done, 1);  // layer 1's accumulator pre-loads (FOLD)
    tc::mbar_init_fence();
  }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_base;
  if (F4) {
    if (warp < 4) tc::tmem_fill16(tmem + ((uint32_t)(warp * 32) << 16) + SF_COL, 0x7f7f7f7fu);  // every scale factor = 1.0 (ue8m0)
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
  }
  const uint32_t idesc = tc::make_idesc_i8(128, 64), idesc_f4 = tc::make_idesc_mxf4(128, 64), sf = tmem + SF_COL;
  const uint32_t sA_u = tc::smem_u32(smem), sB1_u = tc::smem_u32(sB1), sP_u = tc::smem_u32(sP), sB0_u = tc::smem_u32(sB0);
  bool failed = false;
#ifdef BNNB_TRACE
  __shared__ long long tr[7][4][8];  // [event][image 8..11][tile]
#define TR(e, k, t) if (blockIdx.x == 0 && lane == 0 && (k) >= 8 && (k) < 12) tr[e][(k) - 8][t] = clock64()
#else
#define TR(e, k, t)
#endif
  const long long first = blockIdx.x, stride = gridDim.x;
  const int K = first < n_img ? (int)((n_img - first + stride - 1) / stride) : 0;  // images of this CTA

  // window plane of image kk, built by the helper warps two images ahead of layer 0's MMAs.  (Shared-memory reads wait behind the
  // MMAs' operand stream for hundreds of clocks, so this and the pool are kept off the epilogue warps' per-tile chains.)
  auto build_p = [&](int kk) {
    const int hx = tid - 19 * 32;
    const int buf = kk & 1;
    const uint32_t par = (uint32_t)(kk >> 1) & 1;
    if (!failed && !tc::mbar_wait<TFS(0)>(&img_full[buf], par)) failed = true;
    if (kk >= 2 && !failed && !tc::mbar_wait<TFS(1)>(&p_empty[buf], par ^ 1)) failed = true;  // layer-0 MMAs of image kk-2 are done with this plane
    const uint32_t* iw = reinterpret_cast<const uint32_t*>(sImg + (size_t)buf * IMG_BYTES);
    uint8_t* sPk = sP + (size_t)buf * S::P_STRIDE;
#pragma unroll PLANEU
    for (int m = hx; kPlane && m < 1024 + 64; m += NH * 32) {  // rows up to 957 + 64 are read by valid outputs; the rest only need to be defined
      const int b = 3 * m, w = b >> 2, sh = (b & 3) * 8;
      uint32_t x0 = 0, x1 = 0, x2 = 0, x3 = 0, x4 = 0;
      if (w + 4 < IMG_BYTES / 4) { x0 = iw[w]; x1 = iw[w + 1]; x2 = iw[w + 2]; x3 = iw[w + 3]; x4 = iw[w + 4]; }
      *reinterpret_cast<uint4*>(sPk + (size_t)m * 16) =
          make_uint4(__funnelshift_r(x0, x1, sh), __funnelshift_r(x1, x2, sh), __funnelshift_r(x2, x3, sh), __funnelshift_r(x3, x4, sh));
    }
    tc::fence_async_smem();
    tc::mbar_arrive(&p_full[buf]);
    if (warp == 19) { TR(6, kk, 2); }
  };
  if (role == 1) {
    // ================= half B (8 warps): layer-0 epilogue into layer-1 operand buffer k&1; team = 32-channel half ================
    int thrA[32];  // first thresholds in registers: the epilogue's shared-memory traffic competes with the MMA operand reads
#pragma unroll
    for (int c = 0; c < 32; c++)  // |S| <= 27 * 128: thresholds outside [-3457, 3456] decide like the nearest bound
      thrA[c] = CP ? max(min(sthr0[2 * (team * 32 + c)], 3456), -3457) : sthr0[2 * (team * 32 + c)];
    uint32_t D0[16];  // TFOLD: (thr1 - thr0) of channels 2j (low half) and 2j+1 (high half), both clamped like thr0
    if (TFOLD) {
#pragma unroll
      for (int j = 0; j < 16; j++) {
        const int da = max(min(sthr0[2 * (team * 32 + 2 * j) + 1], 3456), -3457) - thrA[2 * j];
        const int db = max(min(sthr0[2 * (team * 32 + 2 * j + 1) + 1], 3456), -3457) - thrA[2 * j + 1];
        D0[j] = (uint32_t)db * 65536u + (uint32_t)da;
      }
    }
    int s0 = 0;
    uint32_t p0 = 0;
    for (int k = 0; k < K; k++) {
      const int buf = k & 1;
      const uint32_t par = (uint32_t)(k >> 1) & 1;
      // the layer-1 MMAs that read this operand buffer two images ago must have completed before it is overwritten
      if (warp == 8) { TR(6, k, 5); }
      if (k >= 2 && !failed && !tc::mbar_wait<TFS(3)>(&abuf_empty[buf], par ^ 1)) failed = true;
      if (warp == 8) { TR(6, k, 4); }
      uint8_t* sA = smem + (size_t)buf * L1_A;
#pragma unroll
      for (int t = 0; t < L0_TILES; t++) {
        const int slot = s0, spar = p0;  // tile u = 8k + t uses slot u % L0_SLOTS for the (u / L0_SLOTS)-th time
        if (++s0 == L0_SLOTS) { s0 = 0; p0 ^= 1u; }
        if (!failed && !tc::mbar_wait<TFS(4)>(&f0[slot], spar)) failed = true;
        tc::fence_after_sync();
        const int m = t * 128 + hw * 32 + lane, y = m >> 5, x = m & 31;
        const bool valid = x < 30 && y < 30;
        uint32_t r[32];
        if (FOLD && !kEpi) {
        } else if (FOLD) {
          uint32_t h[16];  // h[j]: low 16 bits of channels 2j (low half) and 2j+1 (high half)
          tc::tmem_ld32_pack16(tmem + ((uint32_t)(hw * 32) << 16) + L0_ACC + (uint32_t)slot * 64 + team * 32, h);
          tc::tmem_ld_wait_dep16(h);
          uint32_t v[4];
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const uint32_t lo = sign_bytes(h[4 * q], h[4 * q + 1]), hi = sign_bytes(h[4 * q + 2], h[4 * q + 3]);  // channels 8q+0..3, 8q+4..7
            v[q] = (lo & 0x0F0F0F0Fu) | (hi & 0xF0F0F0F0u);
          }
          if (valid) *reinterpret_cast<uint4*>(sA + (size_t)team * L1_PLANE + (size_t)(y * 30 + x) * 16) = make_uint4(v[0], v[1], v[2], v[3]);
        } else if (TFOLD) {
          uint32_t h[16];  // h[j]: U = thr0 - S + 32768 of channels 2j (low half) and 2j+1 (high half); bit 15 of a half = NOT (thr0 < S)
          tc::tmem_ld32_pack16(tmem + ((uint32_t)(hw * 32) << 16) + L0_ACC + (uint32_t)slot * 64 + team * 32, h);
          tc::tmem_ld_wait_dep16(h);
          uint32_t v[4];
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const uint32_t y0 = h[4 * q] + D0[4 * q], y1 = h[4 * q + 1] + D0[4 * q + 1], y2 = h[4 * q + 2] + D0[4 * q + 2], y3 = h[4 * q + 3] + D0[4 * q + 3];
            // nibble masks (FOLD's nibble order) of NOT t0 and NOT t1
            const uint32_t m0 = (sign_bytes(h[4 * q], h[4 * q + 1]) & 0x0F0F0F0Fu) | (sign_bytes(h[4 * q + 2], h[4 * q + 3]) & 0xF0F0F0F0u);
            const uint32_t m1 = (sign_bytes(y0, y1) & 0x0F0F0F0Fu) | (sign_bytes(y2, y3) & 0xF0F0F0F0u);
            // neither threshold exceeded -> 0xA (-1.0), both -> 0x2 (+1.0), one -> 0x0
            v[q] = (~(m0 ^ m1) & 0xAAAAAAAAu) & (m0 | 0x22222222u);
          }
          if (valid) *reinterpret_cast<uint4*>(sA + (size_t)team * L1_PLANE + (size_t)(y * 30 + x) * 16) = make_uint4(v[0], v[1], v[2], v[3]);
        } else {
        tc::tmem_ld32(tmem + ((uint32_t)(hw * 32) << 16) + L0_ACC + (uint32_t)slot * 64 + team * 32, r);
        tc::tmem_ld_wait_dep(r);
        }
        if (CP) {  // done above
        } else {
#pragma unroll
          for (int c16 = 0; c16 < 2; c16++) {  // one 16-channel chunk = one uint4 store
            uint32_t v[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
              uint32_t wv = 0;
#pragma unroll
              for (int e = 0; e < 4; e++) {
                const int c = c16 * 16 + q * 4 + e, n = team * 32 + c;
                const int sv = (int)r[c];
                int val = thrA[c] < sv;                                           // {0,1}
                if (OUTM == OUT_TERN) val = val + (sthr0[2 * n + 1] < sv) - 1;    // {-1,0,+1}
                wv |= (uint32_t)(val & 0xff) << (8 * e);
              }
              v[q] = wv;
            }
            if (valid) *reinterpret_cast<uint4*>(sA + (size_t)(2 * team + c16) * L1_PLANE + (size_t)(y * 30 + x) * 16) = make_uint4(v[0], v[1], v[2], v[3]);
          }
        }
        if (TAP) {  // parity taps: the MMA sum S = sum w*q of every valid pixel and channel, and the operand row just written
          const long long img = first + (long long)k * stride;
          if (CP) {  // the accumulator holds thr - S (+ 32768) in 32 bits (the epilogue above read only its low halves)
            tc::tmem_ld32(tmem + ((uint32_t)(hw * 32) << 16) + L0_ACC + (uint32_t)slot * 64 + team * 32, r);
            tc::tmem_ld_wait_dep(r);
          }
          if (valid) {
            int32_t* ta = tp.acc0 + ((size_t)img * 900 + (y * 30 + x)) * 64 + team * 32;
#pragma unroll
            for (int c = 0; c < 32; c++) ta[c] = CP ? thrA[c] + (TFOLD ? 32768 : 0) - (int)r[c] : (int)r[c];
            const uint8_t* row = sA + (size_t)(F4 ? team : 2 * team) * L1_PLANE + (size_t)(y * 30 + x) * 16;
            uint4* td = tp.act0 + ((size_t)img * 900 + (y * 30 + x)) * 4 + 2 * team;
            td[0] = *reinterpret_cast<const uint4*>(row);
            td[1] = F4 ? make_uint4(0, 0, 0, 0) : *reinterpret_cast<const uint4*>(row + L1_PLANE);
          }
        }
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&e0[slot]);  // this warp's share of the accumulator has been read (and re-loaded)
        // rows [120t, 120t+120) of the operand planes are final: layer-1 tile t-1 (which reads rows < 128(t-1)+190) may start.
        // (generic-proxy writes, read by the async proxy: fence first)
        tc::fence_async_smem();
        tc::mbar_arrive(&l0_done[buf][t]);
        if (warp == 8) { TR(2, k, t); }
      }
    }
  } else if (role == 2) {
    // ================= issuer + loader (one thread) =============================================================================
    // The tensor pipe is one in-order queue, so the order of issue is the schedule.  Stage s interleaves layer 1 of image s with
    // layer 0 of image s+1 (two L0 tiles, then one L1 tile): layer 0's MMAs never queue behind more than one layer-1 tile (nine
    // MMAs), and image s+1 is staged well before layer 1 of image s has been issued.
    {  // the whole warp runs the schedule (uniform control flow and addresses); one elected lane issues
      const bool leader = tc::elect_one();
      // descriptor words: every operand here has SBO = 128; the start field (16-byte units) is the only part that moves
      const uint32_t d_hi = tc::desc_hi(128);
      const uint32_t p_lo[2] = {tc::desc_lo(sP_u, 32 * 16), tc::desc_lo(sP_u + S::P_STRIDE, 32 * 16)};
      // FOLD: the second MMA's chunks are ky = 2 and the constant plane one plane further on (LBO = P_BYTES)
      const uint32_t p2_lo[2] = {tc::desc_lo(sP_u, CP ? P_BYTES : 32 * 16), tc::desc_lo(sP_u + S::P_STRIDE, CP ? P_BYTES : 32 * 16)};
      const uint32_t a1_lo[2] = {tc::desc_lo(sA_u, L1_PLANE), tc::desc_lo(sA_u + L1_A, L1_PLANE)};
      const uint32_t b0_lo = tc::desc_lo(sB0_u, 64 * 16), b1_lo = tc::desc_lo(sB1_u, 64 * 16);
      if (warp == 18) tc::griddep_wait();  // the images are the kernel below's output (programmatic dependent launch)
      for (int k = 0; k < 2 && k < K; k++)
        if (leader && warp == 18) {
          tc::mbar_expect_tx(&img_full[k], 3072);
          tc::bulk_g2s(sImg + (size_t)k * IMG_BYTES, in + (first + (long long)k * stride) * 3072, 3072, &img_full[k]);
        }
      if (CP && !failed && !tc::mbar_wait<TFS(5)>(&pre_done, 0)) failed = true;  // accumulators hold their pre-loads
      int s0 = 0;
      uint32_t p0 = 0;
      auto issue_l0 = [&](int k, int t) {
        const int buf = k & 1;
        if (t == 0) {
          TR(6, k, 0);
          if (!failed && !tc::mbar_wait<TFS(5)>(&p_full[buf], (uint32_t)(k >> 1) & 1)) failed = true;
          TR(6, k, 1);
          if (k + 2 < K && leader) {  // image buffer `buf` has been consumed: prefetch image k+2 into it
            tc::mbar_expect_tx(&img_full[buf], 3072);
            tc::bulk_g2s(sImg + (size_t)buf * IMG_BYTES, in + (first + (long long)(k + 2) * stride) * 3072, 3072, &img_full[buf]);
          }
        }
        const int slot = s0;  // tile u = 8k + t uses slot u % L0_SLOTS for the (u / L0_SLOTS)-th time
        const uint32_t spar = p0;
        if (++s0 == L0_SLOTS) { s0 = 0; p0 ^= 1u; }
        TR(0, k, t);
        if (k * L0_TILES + t >= L0_SLOTS && !failed && !tc::mbar_wait<TFS(6)>(&e0[slot], spar ^ 1u)) failed = true;
        tc::fence_after_sync();
        const uint32_t acc = tmem + L0_ACC + (uint32_t)slot * 64;
        const uint32_t a0 = p_lo[buf] + (uint32_t)(t * 128);  // descriptor start fields are in 16-byte units: one row = 1
        if (leader) {
          if (kMma0) tc::mma_i8_lh(acc, a0, d_hi, b0_lo, d_hi, idesc, 0u);                                              // ky 0, 1
          if (kMma0) tc::mma_i8_lh(acc, p2_lo[buf] + (uint32_t)(t * 128 + 64), d_hi, b0_lo + 2 * 64, d_hi, idesc, 1u);  // ky 2, zero / threshold chunk
          tc::commit(&f0[slot]);
          TR(1, k, t);
          if (t == L0_TILES - 1) tc::commit(&p_empty[buf]);
        }
        __syncwarp();
      };
      auto issue_l1 = [&](int k, int t) {
        const int buf = k & 1;
        TR(3, k, t);
        const uint32_t sa_u = sA_u + (uint32_t)buf * L1_A;
        const uint32_t a_lo = a1_lo[buf] + (uint32_t)(t * 128);
        const int g = k * L1_TILES + t, slot = g & 3, u = g >> 2;
        if (u >= 1 && !failed && !tc::mbar_wait<TFS(7)>(&empty[slot], (uint32_t)(u - 1) & 1)) failed = true;
        tc::fence_after_sync();
        const uint32_t acc = tmem + (uint32_t)slot * 64;
        bool firstmma = true;
        if (leader && F4) {
          if (kMma1) tc::mma9_mxf4<1, 30, 2 * 64>(acc, a_lo, b1_lo, d_hi, idesc_f4, sf, CP ? 1u : 0u);
          tc::commit(&full[slot]);
          TR(4, k, t);
        } else if (leader) {
#pragma unroll
        for (int tap = 0; tap < 9; tap++) {
          const uint32_t off = (uint32_t)((tap / 3) * 30 + tap % 3 + t * 128) * 16;
          if (F4) {
            tc::mma_mxf4_lh(acc, a_lo + (uint32_t)((tap / 3) * 30 + tap % 3), d_hi, b1_lo + (uint32_t)(tap * 2 * 64), d_hi, idesc_f4, sf, sf,
                            (CP || !firstmma) ? 1u : 0u);
            firstmma = false;
          } else {
#pragma unroll
            for (int s2 = 0; s2 < 2; s2++) {
              tc::mma_i8(acc, tc::make_desc(sa_u + (uint32_t)(2 * s2) * L1_PLANE + off, L1_PLANE, 128),
                         tc::make_desc(sB1_u + (uint32_t)((tap * 4 + 2 * s2) * 64) * 16, 64 * 16, 128), idesc, firstmma ? 0u : 1u);
              firstmma = false;
            }
          }
        }
        tc::commit(&full[slot]);
        }
        __syncwarp();
      };
      // Three issuer warps.  A lone thread retires about one instruction per 6-7 clocks, so every tile costs its issuer a fixed
      // few hundred clocks (barrier poll, fence, descriptor set-up, commit) on top of ~30 per MMA: measured, the kernel with all MMAs
      // and epilogues removed still took 4 240 clocks per image with one issuer and 3 470 with two.  Warps 16 and 17 issue layer 1
      // (tiles alternate by global tile parity: each warp owns two of the four accumulator slots), warp 18 issues layer 0 and
      // prefetches images.  Layer 1's issue rate stays below the pipe's execution rate, so the in-order MMA queue is short and
      // layer 0's tiles do not wait behind it.
      if (warp == 18) {
        for (int k = 0; k < K; k++) {
#pragma unroll
          for (int t = 0; t < L0_TILES; t++) issue_l0(k, t);
        }
      } else {
        const int w = warp & 1;
        for (int k = 0; k < K; k++) {
          const int buf = k & 1;
          for (int t = 0; t < L1_TILES; t++)
            if (((k * L1_TILES + t) & 1) == w) {
              // layer-1 tile t reads operand rows [128t, 128t+190): complete once layer-0 tile t+1 has been written
              if (!failed && !tc::mbar_wait<TFS(8)>(&l0_done[buf][t + 1], (uint32_t)(k >> 1) & 1)) failed = true;
              issue_l1(k, t);
            }
          if (leader) tc::commit(&abuf_empty[buf]);  // this warp's MMAs of the image; the barrier counts both issuers
          __syncwarp();
        }
      }
    }
  } else if (role == 0) {
    // ================= half A (8 warps): layer-1 epilogue into bit rows; team = 32-channel half ===================================
    constexpr int WPP = 2 * S::OPL;
    using acc_t = typename std::conditional<F4, float, int>::type;  // kind::mxf4 accumulates in fp32 (exact: |acc| <= 1152)
    const acc_t* sthr1t = reinterpret_cast<const acc_t*>(sthr1);
    acc_t thrA[32];
#pragma unroll
    for (int c = 0; c < 32; c++) {
      const acc_t t = sthr1t[2 * (CP ? tfold_channel(team * 32 + c) : team * 32 + c)];
      thrA[c] = t;
      // FOLD pre-load: 1.5 * 2^23 + 32768 + 6 * thr + 5.  Layer 0 writes a set decision as the code 0xF = -6.0, so the MMAs leave
      // U = 6 * thr + 5 - 6 * S + 32768 in the low 16 bits: bit 15 clear <=> 6 * S > 6 * thr + 5 <=> thr < S  (|S| <= 576: thresholds
      // outside [-577, 576] decide like the nearest bound, and the "no threshold" sentinel never fires)
      if (FOLD) thrA[c] = (acc_t)(12582912.0f + 32768.0f + 5.0f + 6.0f * fminf(fmaxf((float)t, -577.0f), 576.0f));
      // TFOLD pre-load: 1.5 * 2^23 + 32768 + thr0 (|S| <= 2 * 576: thresholds outside [-1153, 1152] decide like the nearest bound)
      if (TFOLD) thrA[c] = (acc_t)(12582912.0f + 32768.0f + fminf(fmaxf((float)t, -1153.0f), 1152.0f));
    }
    uint32_t D1[16];  // TFOLD: (thr1 - thr0) of columns 2j (low half) and 2j+1 (high half)
    if (TFOLD) {
#pragma unroll
      for (int j = 0; j < 16; j++) {
        const float a0 = (float)thrA[2 * j] - (12582912.0f + 32768.0f), b0 = (float)thrA[2 * j + 1] - (12582912.0f + 32768.0f);
        const float a1 = fminf(fmaxf((float)sthr1t[2 * tfold_channel(team * 32 + 2 * j) + 1], -1153.0f), 1152.0f);
        const float b1 = fminf(fmaxf((float)sthr1t[2 * tfold_channel(team * 32 + 2 * j + 1) + 1], -1153.0f), 1152.0f);
        D1[j] = (uint32_t)(int)(b1 - b0) * 65536u + (uint32_t)(int)(a1 - a0);
      }
    }
    auto preload1 = [&](int slot) {
      tc::tmem_st32(tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * 64 + team * 32, reinterpret_cast<const uint32_t(&)[32]>(thrA[0]));
      tc::tmem_st_wait();
    };
    if (CP) {
#pragma unroll
      for (int slot = 0; slot < 4; slot++) preload1(slot);
      tc::fence_before_sync();
      half_sync(1);
      if (ht == 0) tc::mbar_arrive(&pre_done);
    }
    for (int k = 0; k < K; k++) {
      const int bb = k & (S::NBITS - 1);  // bit-row buffer: the helpers pool image k while this half is on image k+1
      uint32_t* sbk = sbits + (size_t)bb * (S::BITS / 4);
      if (k >= S::NBITS && !failed && !tc::mbar_wait<TFS(9)>(&bits_empty[bb], (uint32_t)(k / S::NBITS - 1) & 1)) failed = true;
      for (int t = 0; t < L1_TILES; t++) {
        const int g = k * L1_TILES + t, slot = g & 3, u = g >> 2;
        if (!failed && !tc::mbar_wait<TFS(10)>(&full[slot], (uint32_t)u & 1)) failed = true;
        tc::fence_after_sync();
        const int m = t * 128 + hw * 32 + lane;
        uint32_t r[32];
        if (!CP || TAP) {
          tc::tmem_ld32(tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * 64 + team * 32, r);
          tc::tmem_ld_wait_dep(r);
        }
        uint32_t b0 = 0, b1 = 0;
        if (FOLD && !kEpi) {
        } else if (TFOLD) {
          uint32_t h[16];  // low 16 bits of the accumulators = U = thr0 - S + 32768, columns 2j / 2j+1 in the halves
          tc::tmem_ld32_pack16(tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * 64 + team * 32, h);
          tc::tmem_ld_wait_dep16(h);
          uint32_t n0 = 0, n1 = 0;  // bit c = NOT (thr0 < S) / NOT (thr1 < S) of channel c
#pragma unroll
          for (int g = 0; g < 8; g++) {  // registers 2g, 2g+1 = channels g, 8+g, 16+g, 24+g -> bit g of the four bytes
            n0 |= sign_bytes(h[2 * g], h[2 * g + 1]) & (0x01010101u << g);
            n1 |= sign_bytes(h[2 * g] + D1[2 * g], h[2 * g + 1] + D1[2 * g + 1]) & (0x01010101u << g);
          }
          b0 = ~(n0 ^ n1);  // not zero: both thresholds exceeded, or neither
          b1 = ~(n0 | n1);  // +1: both
        } else if (FOLD) {
          uint32_t h[16];  // low 16 bits of the accumulators, columns 2j / 2j+1 in the halves: bit 15 = NOT (thr < S)
          tc::tmem_ld32_pack16(tmem + ((uint32_t)(hw * 32) << 16) + (uint32_t)slot * 64 + team * 32, h);
          tc::tmem_ld_wait_dep16(h);
          uint32_t n0 = 0;
#pragma unroll
          for (int g = 0; g < 8; g++) n0 |= sign_bytes(h[2 * g], h[2 * g + 1]) & (0x01010101u << g);  // channels g, 8+g, 16+g, 24+g
          b0 = ~n0;
        } else {
#pragma unroll
          for (int c = 0; c < 32; c++) {
            const acc_t sv = F4 ? (acc_t)__uint_as_float(r[c]) : (acc_t)(int)r[c];
            const int t0 = thrA[c] < sv;
            if (OUTM == OUT_BIT) b0 |= (uint32_t)t0 << c;
            else { const int t1 = sthr1t[2 * (team * 32 + c) + 1] < sv; b0 |= (uint32_t)(1 ^ t0 ^ t1) << c; b1 |= (uint32_t)(t0 & t1) << c; }
          }
        }
        if (kEpi || !FOLD) sbk[(size_t)m * WPP + team] = b0;
        if (OUTM == OUT_TERN) sbk[(size_t)m * WPP + 2 + team] = b1;
        if (TAP) {  // parity taps: the MMA sum of every valid output pixel and channel (pre-load un-folded)
          const int y = m / 30, x = m - y * 30;
          if (y < 28 && x < 28) {
            int32_t* ta = tp.acc1 + ((size_t)(first + (long long)k * stride) * 784 + (y * 28 + x)) * 64 + team * 32;
#pragma unroll
            for (int c = 0; c < 32; c++) {
              if (TFOLD) ta[tfold_channel(c)] = (int)(__float_as_uint((float)thrA[c]) - r[c]);  // both are 0x4B400000 + integer
              else if (FOLD) ta[tfold_channel(c)] = (int)(__float_as_uint((float)thrA[c]) - r[c]) / 6;       // pre-load - accumulator = 6 * S
              else ta[c] = F4 ? __float2int_rn(__uint_as_float(r[c])) : (int)r[c];
            }
          }
        }
        if (CP && kEpi) preload1(slot);
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&empty[slot]);  // this warp's share of the accumulator has been read (and re-loaded)
        if (warp == 0) { TR(5, k, t); }
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&bits_full[bb]);  // release: this warp's bit rows of image k are written
    }
  } else {
    // ================= helpers (NH warps): window planes two images ahead; 2x2 max-pool of the bit rows and store =================
    constexpr int WPP = 2 * S::OPL;
    const int hx = tid - 19 * 32;
    if (K > 0) build_p(0);
    if (K > 1) build_p(1);
    for (int k = 0; k < K; k++) {
      const long long img = first + (long long)k * stride;
      if (k + 2 < K) build_p(k + 2);
      const int bb = k & (S::NBITS - 1);
      const uint32_t* sbk = sbits + (size_t)bb * (S::BITS / 4);
      if (!failed && !tc::mbar_wait<TFS(2)>(&bits_full[bb], (uint32_t)(k / S::NBITS) & 1)) failed = true;
      // 2x2 max-pool over the 28x28 valid outputs (row index y*30 + x) -> 14x14, bit-packed
#pragma unroll POOLU
      for (int i = hx; kPool && i < 196 * 2; i += NH * 32) {
        const int j = i & 1, op = i >> 1, oy = op / 14, ox = op % 14;
        const uint32_t* r00 = sbk + (size_t)((2 * oy) * 30 + 2 * ox) * WPP;
        const uint32_t *r01 = r00 + WPP, *r10 = r00 + 30 * WPP, *r11 = r10 + WPP;
        uint32_t* dst = out + ((size_t)img * 196 + op) * WPP;
        if (OUTM == OUT_BIT) {
          const uint32_t w = r00[j] | r01[j] | r10[j] | r11[j];
          if (out_fmt == ACT_F4P) reinterpret_cast<uint4*>(out)[((size_t)img * 2 + j) * 196 + op] = f4_from_bits(w);  // [img][plane][pixel]
          else dst[j] = w;
        } else {
          const uint32_t P = r00[2 + j] | r01[2 + j] | r10[2 + j] | r11[2 + j];
          const uint32_t M = (r00[j] & ~r00[2 + j]) & (r01[j] & ~r01[2 + j]) & (r10[j] & ~r10[2 + j]) & (r11[j] & ~r11[2 + j]);
          if (out_fmt == ACT_F4P) reinterpret_cast<uint4*>(out)[((size_t)img * 2 + j) * 196 + op] = f4_from_tern(P | M, P);  // [img][plane][pixel]
          else { dst[j] = P | M; dst[2 + j] = P; }
        }
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&bits_empty[bb]);
      if (warp == 19) { TR(6, k, 3); }
    }
  }
  if (failed && status) *status = 1;
  tc::fence_before_sync();
  __syncthreads();
#ifdef BNNB_TRACE
  if (blockIdx.x == 0 && tid == 0 && K >= 12) {
    const long long t0 = tr[0][0][0];
    for (int e = 0; e < 7; e++)
      for (int k = 0; k < 4; k++)
        printf("TRACE e%d k%d: %lld %lld %lld %lld %lld %lld %lld %lld\n", e, k + 8, tr[e][k][0] - t0, tr[e][k][1] - t0, tr[e][k][2] - t0, tr[e][k][3] - t0,
               tr[e][k][4] - t0, tr[e][k][5] - t0, tr[e][k][6] - t0, (e < 3 || e == 6) ? tr[e][k][7] - t0 : 0);
  }
#endif
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

}  // namespace

template <int OUTM, bool F4>
static void launch01(const DevLayer& L0, const DevLayer& L1, const int8_t* in, uint32_t* out, int n_img, int grid, int out_fmt, int* status, const Tap01& tp,
                     cudaStream_t s) {
  using S = Smem01<OUTM, F4>;
  static_assert(S::TOTAL <= 226 * 1024, "shared memory budget");
  if (tp.acc0) {  // parity taps: the TAP instantiation of the same kernel
    cudaFuncSetAttribute(k_cnv01_tc<OUTM, F4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL);
    k_cnv01_tc<OUTM, F4, true><<<grid, NT01, S::TOTAL, s>>>(L0, L1, in, out, n_img, out_fmt, status, tp);
    return;
  }
  cudaFuncSetAttribute(k_cnv01_tc<OUTM, F4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL);
  launch_maybe_pdl(k_cnv01_tc<OUTM, F4, false>, grid, NT01, S::TOTAL, s, L0, L1, in, out, n_img, out_fmt, status, tp);
}

bool launch_cnv01_tc(const DevLayer& L0, const DevLayer& L1, const int8_t* in, uint32_t* out, int n_img, int num_sms, bool fp4, int out_fmt,
                     int* status, const Tap01& tp, cudaStream_t s) {
  if (out_fmt != ACT_BITS && out_fmt != ACT_F4P) return false;
  if (n_img <= 0 || !L1.w8 || !L1.w4 || !L1.thr_tc || L0.d.N != 64 || L1.d.N != 64 || L1.d.ifm_dim != 30) return false;
  const int grid = n_img < num_sms ? n_img : num_sms;
  const bool bit = L1.d.out_mode == OUT_BIT;
  if (bit && fp4) launch01<OUT_BIT, true>(L0, L1, in, out, n_img, grid, out_fmt, status, tp, s);
  else if (bit) launch01<OUT_BIT, false>(L0, L1, in, out, n_img, grid, out_fmt, status, tp, s);
  else if (fp4) launch01<OUT_TERN, true>(L0, L1, in, out, n_img, grid, out_fmt, status, tp, s);
  else launch01<OUT_TERN, false>(L0, L1, in, out, n_img, grid, out_fmt, status, tp, s);
  return true;
}

}  // namespace bnnb
