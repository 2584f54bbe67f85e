"""tcgen05 kind::i8 descriptor probe (bnnb_tc_probe): the operand layout and the shifted-start implicit im2col that the
tensor-core conv kernels rely on, checked against numpy int32 matmul.  Bit-exact."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,K,shift", [(64, 32, 0), (64, 64, 0), (64, 576, 0), (128, 128, 0), (256, 64, 0), (64, 64, 1), (64, 64, 31), (128, 96, 62), (32, 32, 3)])
def test_probe_matches_numpy(N, K, shift):
    from bnn import native
    rng = np.random.default_rng(N * 1000 + K + shift)
    rows = 128 + ((shift + 7) // 8) * 8
    A = rng.integers(-128, 128, (rows, K), dtype=np.int8)
    B = rng.integers(-2, 2, (N, K), dtype=np.int8)
    D, _ = native.tc_probe(A, B, shift)
    want = A[shift:shift + 128].astype(np.int32) @ B.astype(np.int32).T
    assert np.array_equal(D, want)


def test_probe_rate_report():
    from bnn import native
    rng = np.random.default_rng(0)
    out = {}
    for N in (64, 128, 256):
        K = 256
        A = rng.integers(-1, 2, (128, K), dtype=np.int8)
        B = rng.integers(-1, 2, (N, K), dtype=np.int8)
        reps = 512
        _, c1 = native.tc_probe(A, B, 0, 1)
        D, c = native.tc_probe(A, B, 0, reps)
        mmas = reps * K // 32
        out[N] = (c - c1) / (mmas - K // 32)
        assert np.array_equal(D, reps * (A.astype(np.int32) @ B.astype(np.int32).T))
    print("tcgen05 kind::i8 M=128 K=32 cycles per MMA by N:", out, "MAC/clk/SM:", {n: 128 * n * 32 / v for n, v in out.items()})


@pytest.mark.parametrize("N,K,shift", [(64, 64, 0), (64, 576, 0), (128, 128, 5), (256, 128, 0), (64, 128, 31), (32, 64, 2)])
def test_fp4_probe_matches_numpy(N, K, shift):
    """kind::mxf4 with unit scales as an exact small-integer MMA ({-1,0,1} activations x {-2..1} weights)."""
    from bnn import native
    rng = np.random.default_rng(N * 7 + K + shift)
    rows = 128 + ((shift + 7) // 8) * 8
    A = rng.integers(-1, 2, (rows, K))
    B = rng.integers(-2, 2, (N, K))
    D, _ = native.tc_probe_fp4(A, B, shift)
    want = A[shift:shift + 128].astype(np.int32) @ B.astype(np.int32).T
    assert np.array_equal(D, want)


def test_fp4_probe_rate_report():
    from bnn import native
    rng = np.random.default_rng(1)
    out = {}
    for N in (64, 128, 256):
        K = 512
        A = rng.integers(0, 2, (128, K))
        B = rng.integers(-1, 2, (N, K))
        reps = 256
        _, c1 = native.tc_probe_fp4(A, B, 0, 1)
        D, c = native.tc_probe_fp4(A, B, 0, reps)
        mmas = reps * K // 64
        out[N] = (c - c1) / (mmas - K // 64)
        assert np.array_equal(D, reps * (A.astype(np.int32) @ B.astype(np.int32).T))
    print("tcgen05 kind::mxf4 M=128 K=64 cycles per MMA by N:", out, "MAC/clk/SM:", {n: 128 * n * 64 / v for n, v in out.items()})


@pytest.mark.parametrize("N,K", [(64, 576), (128, 1152), (32, 2304)])
def test_fp4_accumulator_preload_is_exact(N, K):
    """Threshold folding: the accumulator is pre-loaded with -(thr + 1/2) and the MMAs accumulate on top; the sign bit of the
    result is then the (negated) threshold decision.  Checks that fp32 accumulation stays exact with pre-loads as large as a
    16-bit threshold field (+-32768.5) and for the 3e38 'never fires' sentinel."""
    from bnn import native
    rng = np.random.default_rng(N + K)
    A = rng.integers(-1, 2, (128, K))
    B = rng.integers(-2, 2, (N, K))
    thr = rng.integers(-600, 600, N).astype(np.float64)
    thr[:8] = [32767, -32768, 65535, -65536, 0, -1, 1, 40000]
    bias = (-(thr + 0.5)).astype(np.float32)
    bias[8] = -3.0e38
    D, _ = native.tc_probe_fp4(A, B, 0, 1, bias=bias)
    S = (A.astype(np.int64) @ B.astype(np.int64).T).astype(np.float64)
    want = S + bias.astype(np.float64)[None, :]
    ok = np.ones(N, bool); ok[8] = False
    assert np.array_equal(D[:, ok].astype(np.float64), want[:, ok])
    assert np.all(D[:, 8] < -1e38)
    # the decision itself
    assert np.array_equal(~np.signbit(D[:, ok]), S[:, ok] > thr[None, ok])


@pytest.mark.parametrize("N,K", [(64, 576), (128, 1152), (32, 2304)])
def test_fp4_biased_preload_keeps_the_integer_in_the_low_mantissa_bits(N, K):
    """The packed 16-bit epilogues (f4_common.cuh): the accumulator is pre-loaded with 1.5 * 2^23 + 32767 - thr, the MMAs add the exact
    sum S, and the low 16 bits of the fp32 PATTERN are then the integer S - thr - 1 + 32768, whose bit 15 is the decision thr < S.
    That needs the tensor core's fp32 accumulation to be exact down to the last mantissa bit at this magnitude: checked here for
    +-1 / +-2 operands over the K of the real layers, and once more at 1.5 * 2^24 (ulp 2) with even products."""
    from bnn import native
    rng = np.random.default_rng(7 * N + K)
    A = rng.integers(-1, 2, (128, K))
    B = rng.integers(-2, 2, (N, K))
    smax = 2 * K
    thr = np.clip(rng.integers(-smax - 400, smax + 400, N), -smax - 1, smax).astype(np.int64)
    bias = (12582912.0 + 32767.0 - thr).astype(np.float32)
    assert np.array_equal(bias.astype(np.int64), 12582912 + 32767 - thr)          # the pre-loads themselves are exact floats
    D, _ = native.tc_probe_fp4(A, B, 0, 1, bias=bias)
    S = A.astype(np.int64) @ B.astype(np.int64).T
    assert np.array_equal(D.astype(np.float64), (S + bias.astype(np.int64)[None, :]).astype(np.float64))
    low = np.ascontiguousarray(D, np.float32).view(np.uint32) & 0xFFFF
    assert np.array_equal(low, (S - thr[None, :] - 1 + 32768) & 0xFFFF)
    assert np.array_equal((low >> 15) & 1, (thr[None, :] < S).astype(np.uint32))
    # ulp 2: even products only, the accumulator's last mantissa bit still takes part
    bias2 = (25165824.0 + 2.0 * (32767 - thr)).astype(np.float32)
    D2, _ = native.tc_probe_fp4(A, 2 * (B // 2), 0, 1, bias=bias2)
    S2 = A.astype(np.int64) @ (2 * (B // 2)).astype(np.int64).T
    assert np.array_equal(D2.astype(np.float64), (S2 + bias2.astype(np.int64)[None, :]).astype(np.float64))


def test_mma_kind_switch_cost_report():
    """Timing report: cost of interleaving kind::i8 MMAs among kind::mxf4 MMAs on the same tensor pipe.
    mix = g > 0: groups of g fp4 MMAs + one int8 MMA; mix = -g: the control, g + 1 fp4 MMAs per group."""
    from bnn import native
    rng = np.random.default_rng(3)
    K, N, reps = 64, 64, 512
    A = rng.integers(0, 2, (128, K))
    B = rng.integers(-1, 2, (N, K))
    out = {}
    for g in (1, 2, 4, 9):
        _, c_mix = native.tc_probe_fp4(A, B, 0, reps, mix=g)
        _, c_ctl = native.tc_probe_fp4(A, B, 0, reps, mix=-g)
        out[g] = {"clk_per_group_mixed": c_mix / reps, "clk_per_group_fp4_only": c_ctl / reps}
    print("kind switch:", out)


def test_tmem_pack16_semantics():
    """The fused kernel's layer-0 epilogue reads its int32 accumulators with tcgen05.ld .pack::16b and takes decisions from bits 15
    and 31 of each register: register j must hold the LOW 16 bits of column 2j (low half) and of column 2j+1 (high half)."""
    from bnn import native
    rng = np.random.default_rng(5)
    cols = rng.integers(0, 2**32, (32, 32), dtype=np.uint64).astype(np.uint32)
    cols[0, :8] = np.array([0, 1, 0xFFFF, 0x8000, 0x7FFF, 0xFFFF8000, 0x00017FFF, 0xFFFFFFFF], np.uint32)
    got = native.tmem_pack_probe(cols)
    lo = cols & 0xFFFF
    assert np.array_equal(got, (lo[:, 0::2] | (lo[:, 1::2] << 16)).astype(np.uint32))
    # a folded layer-0 accumulator (|thr - S| <= 6 913) keeps its sign in bit 15 of the low half
    acc = rng.integers(-6913, 6914, (32, 32)).astype(np.int32)
    got = native.tmem_pack_probe(acc.view(np.uint32))
    assert np.array_equal((got >> 15) & 1, (acc[:, 0::2] < 0).astype(np.uint32))
    assert np.array_equal(got >> 31, (acc[:, 1::2] < 0).astype(np.uint32))


def test_alu_pipe_rate_report():
    """Timing report: the epilogues are one ALU instruction per accumulator column; SHF / PRMT / LOP3 share one pipe (64 lanes per
    clock per SM), integer multiply-add runs beside it on the FMA pipe."""
    from bnn import native
    per_clk = {n: round(native.alu_rate(w) / (148 * 1.965e9), 1) for w, n in ((2, "lop3"), (3, "shf.l.wrap"), (4, "prmt"), (5, "mad.hi"), (6, "mad.lo"))}
    print("thread-instructions / clk / SM:", per_clk)
    assert per_clk["shf.l.wrap"] > 32


def test_tmem_rate_report():
    """Timing report: TMEM -> register (tcgen05.ld) and register -> TMEM (tcgen05.st) bytes per clock per SM by number of warps."""
    from bnn import native
    out = {"ld": {w: round(native.tmem_rate(w, False), 1) for w in (1, 2, 4, 8)}, "st": {w: round(native.tmem_rate(w, True), 1) for w in (1, 4, 8)}}
    print("TMEM bytes/clk/SM:", out)
    assert out["ld"][4] > 0


def test_fp4_shifted_descriptor_rate_report():
    """Timing report: MMA rate when the A descriptor starts at a 16-byte but not 128-byte aligned row (the implicit-im2col taps)."""
    from bnn import native
    rng = np.random.default_rng(5)
    out = {}
    for N in (64, 128):
        K = 512
        B = rng.integers(-1, 2, (N, K))
        for shift in (0, 1, 2, 4, 7, 8):
            A = rng.integers(0, 2, (136, K))
            reps = 256
            _, c1 = native.tc_probe_fp4(A, B, shift, 1)
            D, c = native.tc_probe_fp4(A, B, shift, reps)
            out[(N, shift)] = round((c - c1) / ((reps - 1) * K // 64), 1)
            assert np.array_equal(D, reps * (A[shift:shift + 128].astype(np.int32) @ B.astype(np.int32).T))
    print("clk per kind::mxf4 MMA (M=128, K=64) by (N, row shift):", out)


def test_fp4_lean_issue_rate_report():
    """Timing report: tensor-pipe clocks per MMA when the issuing thread is not the limit (unrolled, one add per descriptor)."""
    from bnn import native
    out = {(N, taps): round(native.tc_rate_fp4(N, taps), 1) for N in (32, 64, 128) for taps in (False, True)}
    print("lean clk per kind::mxf4 MMA (M=128, K=64) by (N, 3x3 tap shifts):", out)
    assert all(v > 0 for v in out.values())


def test_mbarrier_roundtrip_report():
    """Timing report: latency of the producer/consumer handshakes the warp-specialised kernels are built from."""
    from bnn import native
    out = {"arrive/wait round trip clk": round(native.mbar_roundtrip(0), 1), "commit/wait round trip clk": round(native.mbar_roundtrip(1), 1)}
    print("mbarrier ping-pong:", out)
    assert out["arrive/wait round trip clk"] > 0
