// Minimal hand-written tcgen05 / TMEM / mbarrier helpers for sm_100a (inline PTX; no CUTLASS).
// Operand layout used throughout: K-major, no swizzle ("interleave"): a core matrix is 8 rows x 16 bytes stored as 128
// contiguous bytes; SBO = byte distance between 8-row groups, LBO = byte distance between the two 16-byte K chunks of one
// K=32 (int8) instruction.  Descriptor bit layout: version 1 (Blackwell), start/LBO/SBO in 16-byte units.
#pragma once
#include <cstdint>

namespace bnnb {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// shared-memory matrix descriptor, K-major, SWIZZLE_NONE
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version for sm_100
  return d;                // base_offset 0, lbo_mode 0, layout_type 0 (no swizzle)
}

// instruction descriptor for kind::i8: D = int32, A/B = int8 (signed unless a_unsigned), both K-major, dense
__host__ __device__ constexpr uint32_t make_idesc_i8(int M, int N, bool a_signed = true, bool b_signed = true) {
  return (2u << 4) | ((a_signed ? 1u : 0u) << 7) | ((b_signed ? 1u : 0u) << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// arrives (count 1) on an mbarrier when all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {  // one full warp; ncols power of 2 >= 32
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // the allocating warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// make generic-proxy shared-memory writes visible to the async proxy (tcgen05.mma operand reads)
// programmatic dependent launch: a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization starts while its
// predecessor in the stream still runs; griddep_wait blocks until the predecessor has completed and its writes are visible
// (no-ops in an ordinary launch)
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// ---- CTA pairs (cta_group::2): two CTAs of a cluster, one per SM of a TPC, run ONE MMA of M = 256: each keeps its own 128 operand
// rows and accumulator lanes and HALF of the B rows; the leader (cluster rank 0) issues, commits are multicast to both ----------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `cta` of the cluster.  (Default semantics, release at CTA scope: a
// cluster-scope release would wait for the warp's outstanding global stores - measured +50 % on the paired conv kernels.)
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t"
      ".reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t"
      "}\n" ::"r"(smem_u32(bar)), "r"(cta)
      : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void commit_pair(uint64_t* bar) {  // both CTAs' barrier at this offset
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)), "h"((uint16_t)3)
               : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_dst, uint32_t ncols) {  // the same warp of BOTH CTAs, same smem_dst offset
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
#if defined(BNNB_TESTWAIT)
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
#elif defined(BNNB_WAIT_HINT)
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
#else
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
#endif
      "selp.u32 %0, 1, 0, p;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity), "r"(0x989680u)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// 4-D tiled TMA load (cuTensorMapEncodeTiled descriptor in kernel parameter space): box at coordinates (c0, c1, c2, c3) -> shared memory
// in box order (dimension 0 fastest), bytes counted on the mbarrier
__device__ __forceinline__ void tma_load_4d(void* dst_smem, const void* tmap, int c0, int c1, int c2, int c3, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst_smem)),
               "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
// bounded spin: returns false once the wait has lasted BNNB_WAIT_NS of wall time (%globaltimer; legitimate waits are microseconds,
// also under profiler replay or time-slicing, where a poll count says nothing) so that a protocol bug cannot hang the GPU box.
// The caller stops waiting after the first failure and raises the device status flag; the engine turns that into an error.
#ifndef BNNB_WAIT_NS
#define BNNB_WAIT_NS 2000000000ull
#endif
// TF (test first): one non-blocking mbarrier.test_wait before the try_wait loop.  A try_wait whose phase has already completed
// still takes ~350 clocks to return; the test answers at once.  Pays where waits mostly find their barrier complete (the 1-bit fused
// kernel 1.237 -> 1.18 ms, conv_f4 layer 3 0.370 -> 0.359), costs where they mostly do not (2-bit fused kernel 1.53 -> 1.57,
// gemm_f4 +5 %), so each kernel chooses.
template <bool TF = false>
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
  // eight polls per trip: the trip counter costs two ALU-pipe instructions, and the epilogue warps share that pipe
#ifndef BNNB_POLLU
#define BNNB_POLLU 8
#endif
#ifndef BNNB_TF_POLLS
#define BNNB_TF_POLLS 1
#endif
#pragma unroll
  for (int tp = 0; tp < (TF ? BNNB_TF_POLLS : 0); tp++) {  // fast path: a barrier that has already completed its phase answers a non-blocking test at once
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    if (ok) return true;
  }
  uint64_t t0 = 0;
  for (;;) {
    for (uint32_t i = 0; i < (1u << 14) / BNNB_POLLU; i++) {
#pragma unroll
      for (int j = 0; j < BNNB_POLLU; j++)
        if (mbar_try_wait(bar, parity)) return true;
    }
    const uint64_t now = global_timer_ns();  // every 2^14 polls (tens of microseconds)
    if (t0 == 0) t0 = now;
    else if (now - t0 > BNNB_WAIT_NS) return false;
  }
}

// the same for a barrier that CTAs of the cluster arrive on (acquire at cluster scope)
__device__ __forceinline__ bool mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint64_t t0 = 0;
  for (;;) {
    for (uint32_t i = 0; i < (1u << 14); i++)
      if (mbar_try_wait_cluster(bar, parity)) return true;
    const uint64_t now = global_timer_ns();
    if (t0 == 0) t0 = now;
    else if (now - t0 > BNNB_WAIT_NS) return false;
  }
}

// TMEM -> registers: this warp's 32 lanes (rows), 32 consecutive 32-bit columns starting at taddr's column
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
// TMEM -> registers with .pack::16b: 32 consecutive columns, the low 16 bits of columns 2j and 2j+1 in register j (low, high half)
__device__ __forceinline__ void tmem_ld32_pack16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait_dep16(uint32_t (&r)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]),
                 "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
               :
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

}  // namespace tc
}  // namespace bnnb

// ---- block-scaled fp4 (kind::mxf4, e2m1 operands packed two per byte, K = 64 per instruction, fp32 accumulate) --------------
// With every scale factor = 1.0 (ue8m0 0x7F) this is an exact small-integer MMA: e2m1 holds {0, +-1, +-2} exactly and the
// fp32 accumulator is exact below 2^24.
namespace bnnb { namespace tc {
__host__ __device__ constexpr uint32_t make_idesc_mxf4(int M, int N) {
  return (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (1u << 23) | ((uint32_t)(M >> 4) << 24);  // a/b = e2m1, scales ue8m0, K=64
}
__device__ __forceinline__ void mma_mxf4(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t sfa, uint32_t sfb, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t"
      "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb)
      : "memory");
}
// registers -> TMEM: this warp's 32 lanes, 16 consecutive columns, all set to v
__device__ __forceinline__ void tmem_fill16(uint32_t taddr, uint32_t v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(taddr), "r"(v)
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// registers -> TMEM: this warp's 32 lanes, 32 consecutive columns (no wait: pair with tmem_st_wait)
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]),
      "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]),
      "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
}}  // namespace bnnb::tc

// ---- TMA bulk copy (1-D, global -> shared) completing on an mbarrier ---------------------------------------------------------
namespace bnnb { namespace tc {
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// dst/src 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
}}  // namespace bnnb::tc

namespace bnnb { namespace tc {
// tcgen05.wait::ld that also names the loaded registers as operands, so that no use of them can be scheduled above the wait
__device__ __forceinline__ void tmem_ld_wait_dep(uint32_t (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]),
                 "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]),
                 "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]),
                 "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}
}}  // namespace bnnb::tc

namespace bnnb { namespace tc {
// one lane of a converged warp (the others fall through); keeps descriptor arithmetic warp-uniform for the compiler
__device__ __forceinline__ bool elect_one() {
  uint32_t p;
  asm volatile(
      "{\n\t"
      ".reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t"
      "}\n"
      : "=r"(p));
  return p != 0;
}
}}  // namespace bnnb::tc

// ---- lean issue path: descriptors as (lo, hi) words so that per-MMA address arithmetic is one 32-bit add ---------------------------
// A single thread issues at roughly one instruction per five clocks, so the instruction count per MMA decides whether the tensor
// pipe is kept fed (measured: tests/test_gpu_tcgen05.py::test_mma_kind_switch_cost_report, ncu source view of the issuer warp).
namespace bnnb { namespace tc {
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr, uint32_t lbo_bytes) { return ((saddr >> 4) & 0x3fff) | (((lbo_bytes >> 4) & 0x3fff) << 16); }
__device__ __forceinline__ uint32_t desc_hi(uint32_t sbo_bytes) { return ((sbo_bytes >> 4) & 0x3fff) | (1u << 14); }  // version 1 at bit 46
__device__ __forceinline__ void mma_mxf4_lh(uint32_t tmem_d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi, uint32_t idesc, uint32_t sfa,
                                            uint32_t sfb, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b64 a, b;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 a, {%1, %2};\n\t"
      "mov.b64 b, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %5, [%7], [%8], p;\n\t"
      "}\n" ::"r"(tmem_d), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb)
      : "memory");
}
__device__ __forceinline__ void mma_i8_lh(uint32_t tmem_d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b64 a, b;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 a, {%1, %2};\n\t"
      "mov.b64 b, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], a, b, %5, p;\n\t"
      "}\n" ::"r"(tmem_d), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc), "r"(accumulate)
      : "memory");
}
}}  // namespace bnnb::tc

// ---- one 3x3 tile = nine kind::mxf4 MMAs in a single asm block ---------------------------------------------------------------------
// The issuing thread is the scarcest resource of these kernels (about one instruction per 5-7 clocks, ncu), so the whole tap loop is
// handed to ptxas as one block: both descriptors keep their high words, the low words advance by immediates.
//   a_lo: descriptor low word of the tile's first operand row; tap (ky,kx) starts ky*YS + kx*XS rows further (16-byte units): XS = 1,
//         YS = image width for rows in image order; XS = G, YS = width*G for G images interleaved pixel by pixel (conv_f4.cu)
//   b_lo: low word of tap 0's weight block; consecutive taps are BSTEP 16-byte units apart
namespace bnnb { namespace tc {
#define BNNB_MMA9_BODY(CG)                                                                                     \
  asm volatile( \
      "{\n\t" \
      ".reg .pred p, q;\n\t" \
      ".reg .b32 al, bl;\n\t" \
      ".reg .b64 a, b;\n\t" \
      "setp.ne.b32 p, %6, 0;\n\t" \
      "setp.eq.b32 q, 0, 0;\n\t" \
      "mov.b64 a, {%1, %3};\n\t" \
      "mov.b64 b, {%2, %3};\n\t" \
      "tcgen05.mma.cta_group::" #CG ".kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], p;\n\t" \
      "add.u32 al, %1, %15;\n\t add.u32 bl, %2, %7;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t" \
      "tcgen05.mma.cta_group::" #CG ".kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t" \
      "add.u32 al, %1, %16;\n\t add.u32 bl, %2, %8;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t" \
      "tcgen05.mma.cta_group::" #CG ".kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t" \
      "add.u32 al, %1, %17;\n\t add.u32 bl, %2, %9;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t" \
      "tcgen05.mma.cta_group::" #CG ".kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t" \
      "add.u32 al, %1, %18;\n\t add.u32 bl, %2, %10;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t" \
      "tcgen05.mma.cta_group::" #CG ".kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t" \
      "add.u32 al, %1, %19;\n\t add.u32 bl, %2, %11;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t" \
      "tcgen05.mma.cta_group::" #CG ".kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t" \
      "add.u32 al, %1, %20;\n\t add.u32 bl, %2, %12;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t" \
      "tcgen05.mma.cta_group::" #CG ".kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t" \
      "add.u32 al, %1, %21;\n\t add.u32 bl, %2, %13;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t" \
      "tcgen05.mma.cta_group::" #CG ".kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t" \
      "add.u32 al, %1, %22;\n\t add.u32 bl, %2, %14;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t" \
      "tcgen05.mma.cta_group::" #CG ".kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t" \
      "}\n" ::"r"(tmem_d), \
      "r"(a_lo), "r"(b_lo), "r"(d_hi), "r"(idesc), "r"(sf), "r"(acc_first), "n"(BSTEP), "n"(2 * BSTEP), "n"(3 * BSTEP), "n"(4 * BSTEP), "n"(5 * BSTEP), \
      "n"(6 * BSTEP), "n"(7 * BSTEP), "n"(8 * BSTEP), "n"(XS), "n"(2 * XS), "n"(YS), "n"(YS + XS), "n"(YS + 2 * XS), "n"(2 * YS), "n"(2 * YS + XS), "n"(2 * YS + 2 * XS) \
      : "memory");
template <int XS, int YS, int BSTEP, int CG = 1>
__device__ __forceinline__ void mma9_mxf4(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t d_hi, uint32_t idesc, uint32_t sf, uint32_t acc_first) {
  if (CG == 2) { BNNB_MMA9_BODY(2); }
  else { BNNB_MMA9_BODY(1); }
#undef BNNB_MMA9_BODY
}
}}  // namespace bnnb::tc

// ---- one ring slot of the K-streamed GEMM = eight kind::mxf4 MMAs in a single asm block -----------------------------------------------
// Same reason as mma9_mxf4: the issuing thread is the scarce resource.  Slab j of the group: A advances by ASTEP 16-byte units (two
// 2 KB chunks of the operand tile), B by BSTEP units (one slab of the ring slot).  acc_first: accumulate flag of the first MMA.
namespace bnnb { namespace tc {
template <int ASTEP, int BSTEP>
__device__ __forceinline__ void mma8_mxf4(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t d_hi, uint32_t idesc, uint32_t sf, uint32_t acc_first) {
  asm volatile(
      "{\n\t"
      ".reg .pred p, q;\n\t"
      ".reg .b32 al, bl;\n\t"
      ".reg .b64 a, b;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "setp.eq.b32 q, 0, 0;\n\t"
      "mov.b64 a, {%1, %3};\n\t"
      "mov.b64 b, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], p;\n\t"
      "add.u32 al, %1, %7;\n\t add.u32 bl, %2, %14;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "add.u32 al, %1, %8;\n\t add.u32 bl, %2, %15;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "add.u32 al, %1, %9;\n\t add.u32 bl, %2, %16;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "add.u32 al, %1, %10;\n\t add.u32 bl, %2, %17;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "add.u32 al, %1, %11;\n\t add.u32 bl, %2, %18;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "add.u32 al, %1, %12;\n\t add.u32 bl, %2, %19;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "add.u32 al, %1, %13;\n\t add.u32 bl, %2, %20;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(d_hi), "r"(idesc), "r"(sf), "r"(acc_first), "n"(ASTEP), "n"(2 * ASTEP), "n"(3 * ASTEP), "n"(4 * ASTEP), "n"(5 * ASTEP),
      "n"(6 * ASTEP), "n"(7 * ASTEP), "n"(BSTEP), "n"(2 * BSTEP), "n"(3 * BSTEP), "n"(4 * BSTEP), "n"(5 * BSTEP), "n"(6 * BSTEP), "n"(7 * BSTEP)
      : "memory");
}
}}  // namespace bnnb::tc

// the same for four and for two slabs (fc_chain.cu: 256-wide layers, and a layer's first channel block, which follows the layer
// below two slabs at a time).  Slab by slab with runtime indices the issuing thread needs ~260 clocks per MMA (six R2UR each).
namespace bnnb { namespace tc {
template <int ASTEP, int BSTEP>
__device__ __forceinline__ void mma4_mxf4(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t d_hi, uint32_t idesc, uint32_t sf, uint32_t acc_first) {
  asm volatile(
      "{\n\t"
      ".reg .pred p, q;\n\t"
      ".reg .b32 al, bl;\n\t"
      ".reg .b64 a, b;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "setp.eq.b32 q, 0, 0;\n\t"
      "mov.b64 a, {%1, %3};\n\t"
      "mov.b64 b, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], p;\n\t"
      "add.u32 al, %1, %7;\n\t add.u32 bl, %2, %10;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "add.u32 al, %1, %8;\n\t add.u32 bl, %2, %11;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "add.u32 al, %1, %9;\n\t add.u32 bl, %2, %12;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(d_hi), "r"(idesc), "r"(sf), "r"(acc_first), "n"(ASTEP), "n"(2 * ASTEP), "n"(3 * ASTEP), "n"(BSTEP), "n"(2 * BSTEP), "n"(3 * BSTEP)
      : "memory");
}
template <int ASTEP, int BSTEP>
__device__ __forceinline__ void mma2_mxf4(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t d_hi, uint32_t idesc, uint32_t sf, uint32_t acc_first) {
  asm volatile(
      "{\n\t"
      ".reg .pred p, q;\n\t"
      ".reg .b32 al, bl;\n\t"
      ".reg .b64 a, b;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "setp.eq.b32 q, 0, 0;\n\t"
      "mov.b64 a, {%1, %3};\n\t"
      "mov.b64 b, {%2, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], p;\n\t"
      "add.u32 al, %1, %7;\n\t add.u32 bl, %2, %8;\n\t mov.b64 a, {al, %3};\n\t mov.b64 b, {bl, %3};\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], a, b, %4, [%5], [%5], q;\n\t"
      "}\n" ::"r"(tmem_d),
      "r"(a_lo), "r"(b_lo), "r"(d_hi), "r"(idesc), "r"(sf), "r"(acc_first), "n"(ASTEP), "n"(BSTEP)
      : "memory");
}
}}  // namespace bnnb::tc
