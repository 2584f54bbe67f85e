"""bnn.finnthesizer (the exporter row of SURVEY 8(f)): trained .npz archive -> <L>-<PE>-weights.bin / -thres.bin.

The reference exporter (bnn/src/training/finnthesizer.py) is Python 2 and needs the `bitstring` package, so it cannot run here.
Parity is pinned on its OUTPUTS instead: the parameter sets the reference ships (copied under tests/golden/params) were written by
that exporter.  For each of the five networks the test unfolds the shipped files into logical matrices with its own code, builds a
training archive whose batch-norm parameters lead to exactly those weights and thresholds (half of the neurons through the
negative-slope branch), runs the exporter and compares every file byte for byte.  The batch-norm folding arithmetic itself is
pinned to the reference's own functions: tests/golden/make_golden_exporter.py executed makeFCBNComplex / makeFCBNComplex_QNN /
makeConvBNComplex / quantize (finnthesizer.py:149-330) here and committed their outputs (test_bn_folding_equals_reference_functions)."""
import os
import sys

import numpy as np
import pytest

from oracle_py import GOLDEN, Oracle, params_dir, synth_cifar, synth_mnist, mnist_idx_bytes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
os.environ.setdefault("BNN_B200_QUIET", "1")

# unpadded shapes (rows = neurons, cols = synapses) and conv geometry of the two topologies
CNV_SHAPES = [(64, 3, 3), (64, 64, 3), (128, 64, 3), (128, 128, 3), (256, 128, 3), (256, 256, 3), (512, 256), (512, 512), (10, 512)]
LFC_SHAPES = [(1024, 784), (1024, 1024), (1024, 1024), (10, 1024)]


def _fth():
    from bnn import finnthesizer
    return finnthesizer


def unfold(pdir, layer, pe, simd, wbits, nthr, rows_p, cols_p):
    """Shipped per-PE files -> (W [rows_p, cols_p] ints, T [rows_p, nthr] ints or None).  Test-local inverse of the folding."""
    sf, nf = cols_p // simd, rows_p // pe
    W = np.zeros((rows_p, cols_p), np.int64)
    T = np.zeros((rows_p, nthr), np.int64) if nthr else None
    for p in range(pe):
        words = np.fromfile(os.path.join(pdir, "%d-%d-weights.bin" % (layer, p)), "<u8")
        assert words.size == nf * sf
        for s in range(simd):
            f = ((words >> np.uint64(s * wbits)) & np.uint64((1 << wbits) - 1)).astype(np.int64)
            if wbits > 1:
                f = np.where(f >= 1 << (wbits - 1), f - (1 << wbits), f)
            W[p::pe, s::simd] = f.reshape(nf, sf)
        if nthr:
            t = np.fromfile(os.path.join(pdir, "%d-%d-thres.bin" % (layer, p)), "<i8")
            assert t.size == nf * nthr
            T[p::pe, :] = t.reshape(nf, nthr)
    return W, T


def archive_for(W, T, conv, popcount, factor, accu_offset, a_bits, wbits, rng):
    """Float 'training' arrays that export to exactly (W, T): unit-magnitude slopes, alternate neurons with negative slope."""
    rows, cols = W.shape
    neg = (np.arange(rows) % 2 == 1) & (T is not None)                      # no batch-norm after the last CNV layer: no flip
    sign = np.where(neg, -1.0, 1.0)
    Wv = (2 * W - 1) if wbits == 1 else W                                   # 1-bit files hold bits: trained weight = +-1
    wt = (Wv * sign[:, None]).astype(np.float32)
    if conv:
        ofm, ifm, k = conv
        weights = wt.reshape(ofm, k, k, ifm).transpose(0, 3, 1, 2).copy()   # columns are (ky, kx, ifm)
    else:
        weights = wt.T.copy()                                               # [ins, outs]
    if T is None:
        return [weights]
    Tl = (T + accu_offset).astype(np.float64)                               # the exporter subtracts the pad offset again
    bias = np.zeros(rows, np.float32)
    beta = np.zeros(rows, np.float32)
    invstd = np.ones(rows, np.float32)
    if a_bits == 1:
        thres = (2 * Tl[:, 0] - cols) if popcount else Tl[:, 0] / factor   # value the exporter must arrive at after an optional flip
        mean = (thres * sign).astype(np.float32)
        gamma = sign.astype(np.float32)
    else:
        d = np.maximum(Tl[:, 1] - Tl[:, 0], 0)
        centre = (Tl[:, 0] + Tl[:, 1]) / 2
        gamma = (sign / np.where(d > 0, d, 1e-6) * factor).astype(np.float32)   # steps of -+0.5 land on T0 / T1 with a quarter margin
        mean = (np.where(neg, -(centre - 0.25), centre + 0.25) / factor).astype(np.float32)
    return [weights, bias, beta, gamma, mean, invstd]


def build_archive(net, pdir):
    cfg = _fth().NETWORKS[net]
    cnv = cfg["kind"] == "cnv"
    shapes = CNV_SHAPES if cnv else LFC_SHAPES
    rng = np.random.default_rng(0)
    arrays = []
    for l, shp in enumerate(shapes):
        pe, simd, wb, ab, ib = cfg["pe"][l], cfg["simd"][l], cfg["w_int"][l], cfg["a_int"][l], cfg["i_int"][l] + cfg["i_fract"][l]
        rows, cols = shp[0], (shp[1] * shp[2] * shp[2] if len(shp) == 3 else shp[1])
        last = l == len(shapes) - 1
        rows_p = ((rows + 63) // 64 * 64) if last else (rows + pe - 1) // pe * pe
        cols_p = (cols + 63) // 64 * 64 if (not cnv and l == 0) else (cols + simd - 1) // simd * simd
        has_thr = not (cnv and last)
        W, T = unfold(pdir, l, pe, simd, wb, (2 if ab == 2 else 1) if has_thr else 0, rows_p, cols_p)
        # padding as the exporter must reproduce it: weight pad 1-bits / zero values, threshold pad 32767 (saturated 65535)
        assert (W[rows:] == (1 if wb == 1 else 0)).all() and (W[:, cols:] == (1 if wb == 1 else 0)).all()
        if has_thr:
            assert (T[rows:] == 32767).all()
        xnor = wb == 1 and ab == 1 and ib == 1
        accu = 0 if (xnor or wb >= 2) else cols_p - cols
        popcount = xnor if not (cnv and l < 6) else (ab == 1 and l > 0)     # conv layers: popcount unless told otherwise (layer 0)
        factor = 256.0 if (cnv and l == 0) else 1.0
        arrays += archive_for(W[:rows, :cols], T[:rows] if has_thr else None, (shp if len(shp) == 3 else None), popcount, factor, accu,
                              ab, wb, rng)
    return {"arr_%d" % i: a for i, a in enumerate(arrays)}


@pytest.mark.parametrize("net", ["cnvW1A1", "cnvW1A2", "cnvW2A2", "lfcW1A1", "lfcW1A2"])
def test_export_reproduces_shipped_parameter_files(net, tmp_path):
    f = _fth()
    pdir = params_dir(net)
    archive = build_archive(net, pdir)
    npz = str(tmp_path / "trained.npz")
    np.savez(npz, *[archive["arr_%d" % i] for i in range(len(archive))])    # positional arrays are stored as arr_0, arr_1, ...
    out = f.export_network(npz, str(tmp_path / "out"), net)
    shipped = sorted(x for x in os.listdir(pdir) if x.endswith(".bin"))
    assert sorted(x for x in os.listdir(out) if x.endswith(".bin")) == shipped
    for name in shipped:
        a, b = open(os.path.join(pdir, name), "rb").read(), open(os.path.join(out, name), "rb").read()
        assert a == b, name
    assert open(os.path.join(out, "classes.txt")).read() == open(os.path.join(pdir, "classes.txt")).read().rstrip("\n")


def test_threshold_arithmetic_by_hand():
    f = _fth()
    # FC, XNOR-popcount: thres = mean - bias - beta/(gamma*invstd) = 10 - 1 - 2/(0.5*2) = 7 -> int((8 + 7)/2) = 7;
    # second neuron has a negative slope: thres = -(3 - 0 - 1/(-1)) = -4 -> int((8 - 4)/2) = 2 and its weights are negated
    w = np.array([[0.3, -0.2]] * 8, np.float32)
    W, T = f.make_fc_bn(w, np.float32([1, 0]), np.float32([2, 1]), np.float32([0.5, -1]), np.float32([10, 3]), np.float32([2, 1]))
    assert T.tolist() == [7, 2] and W.tolist() == [[1] * 8, [1] * 8]
    # signed-sum variant truncates toward zero: thres = -2.5 -> -2
    _, T = f.make_fc_bn(w, np.float32([0, 0]), np.float32([0, 0]), np.float32([1, 1]), np.float32([-2.5, 2.5]), np.float32([1, 1]), use_popcount=False)
    assert T.tolist() == [-2, 2]
    # two-bit activations: steps -0.5 / +0.5; slope 0.25 -> mean +- 2; floor, and ceil of the negation for the flipped neuron
    _, T = f.make_fc_bn_qnn(w, np.float32([0, 0]), np.float32([0, 0]), np.float32([0.25, -0.25]), np.float32([10.3, 10.3]), np.float32([1, 1]), 0, 0, 1, 2)
    assert T.tolist() == [[8, 12], [-12, -8]]
    # convolution, layer-0 style: 1/256 units, no popcount
    cw = np.ones((2, 3, 3, 3), np.float32)
    cw[1] = -1
    W, T = f.make_conv_bn(cw, np.float32([0, 0]), np.float32([0, 0]), np.float32([1, -1]), np.float32([0.36, 0.36]), np.float32([1, 1]), True,
                          0, 0, 7, 1, 1, 1, use_popcount=False, thres_bits=24, thres_int_bits=16)
    assert T.tolist() == [92.0, -92.0] and W.tolist() == [[1] * 27, [1] * 27]     # floor(92.16) / ceil(-92.16); flipped weights
    # popcount conv: fanin 27, thres 5 -> int(32/2) = 16
    _, T = f.make_conv_bn(cw, np.float32([0, 0]), np.float32([0, 0]), np.float32([1, 1]), np.float32([5.9, -5.9]), np.float32([1, 1]), True, 0, 0, 0, 1, 1, 1)
    assert T.tolist() == [16, 10]                                                   # floor(-5.9) = -6 -> int(21/2) = 10
    assert f.quantize(np.float32([-0.0, 0.0, -1e-9]), 1, 0).tolist() == [1, 1, 0]
    assert f.quantize(np.float32([0.5, -0.5, 1.49, -1.5]), 2, 0).tolist() == [1.0, 0.0, 1.0, -1.0]   # round half up, no clipping


def test_channel_interleave_and_column_order():
    f = _fth()
    w = np.arange(2 * 3 * 2 * 2, dtype=np.float32).reshape(2, 3, 2, 2) + 1       # [ofm, ifm, ky, kx], all positive
    ones = np.ones(2, np.float32)
    Wq, _ = f.make_conv_bn(w, 0 * ones, 0 * ones, ones, ones, ones, True, 0, 0, 0, 2, 2, 2)
    assert Wq[0].tolist() == [w[0, c, ky, kx] for ky in range(2) for kx in range(2) for c in range(3)]
    Wq, _ = f.make_conv_bn(w, 0 * ones, 0 * ones, ones, ones, ones, False, 0, 0, 0, 2, 2, 2)
    assert Wq[0].tolist() == w[0].reshape(-1).tolist()
    # first FC layer after a 2-channel conv with 3 pixels per channel: column (chan, pix) moves to (pix, chan)
    r = f.WeightReader({"arr_0": np.arange(6, dtype=np.float32).reshape(6, 1) + 1}, True)
    r.num_interleave_channels = 2
    W, _ = r.read_fc_no_thresholds(0, 0, 0, 2, 1, 2)
    assert W.tolist() == [[1, 4, 2, 5, 3, 6]] and r.num_interleave_channels == 0


def test_proc_elem_mem_folding_and_words():
    f = _fth()
    # 2-bit weights, PE 2, SIMD 4: row r -> PE r % 2; field s of word sf holds column sf*4 + s, two's complement
    W = np.array([[1, -1, 0, -2, 1, 1, 1, 1], [0, 0, 0, 0, -1, -1, -1, -1], [1, 0, 0, 0, 0, 0, 0, 0]], np.int64)
    m = f.ProcElemMem(2, 4, 4, 2, 2, 2, 2)
    m.add_matrix(W, np.array([[1, 2], [3, 40000], [-40000, 6]]), 8, 4)
    assert m.weight_mem.tolist() == [[0b10001101, 0b01010101, 0b00000001, 0], [0, 0b11111111, 0, 0]]
    assert m.threshold_words().tolist() == [[[1, 2], [-32768, 6]], [[3, 32767], [32767, 32767]]]
    # 1-bit weights with a non-XNOR datapath (2-bit activations): pad columns are 1-bits and their count leaves the thresholds
    m = f.ProcElemMem(1, 8, 1, 1, 1, 2, 1)
    m.add_matrix(np.array([[1, 0, 1]]), np.array([[10, 20]]), 8, 1)
    assert m.weight_mem.tolist() == [[0b11111101]] and m.accu_offset == 5 and m.threshold_words().tolist() == [[[5, 15]]]
    with pytest.raises(ValueError):
        f.ProcElemMem(1, 33, 1, 1, 2, 1, 1)                                    # 33 x 2 bits do not fit a 64-bit word


def _random_archive(net, rng):
    """A random 'trained' network of the right shapes (weights uniform, batch-norm statistics of plausible magnitude)."""
    cfg = _fth().NETWORKS[net]
    cnv = cfg["kind"] == "cnv"
    arrays = []
    for l, shp in enumerate(CNV_SHAPES if cnv else LFC_SHAPES):
        rows = shp[0]
        if len(shp) == 3:
            w = rng.uniform(-1, 1, (rows, shp[1], shp[2], shp[2])).astype(np.float32)
            fanin = shp[1] * 9
        else:
            w = rng.uniform(-1, 1, (shp[1], rows)).astype(np.float32)
            fanin = shp[1]
        arrays.append(w if cfg["w_int"][l] == 1 else np.round(w * 1.4).astype(np.float32))
        if cnv and l == 8:
            break
        scale = 0.3 if (cnv and l == 0) else np.sqrt(fanin)
        arrays += [rng.normal(0, 0.1, rows).astype(np.float32), rng.normal(0, 0.5, rows).astype(np.float32),
                   (rng.uniform(0.5, 1.5, rows) * rng.choice([-1, 1], rows, p=[0.2, 0.8])).astype(np.float32),
                   rng.normal(0, 0.3 * scale, rows).astype(np.float32), (rng.uniform(0.5, 1.5, rows) / scale).astype(np.float32)]
    return {"arr_%d" % i: a for i, a in enumerate(arrays)}


@pytest.mark.parametrize("net", ["cnvW1A2", "lfcW1A1"])
def test_exported_random_network_loads_in_the_oracle(net, tmp_path):
    """A freshly 'trained' archive exports to a directory the loader accepts (file sizes, layout) and that is not degenerate."""
    f = _fth()
    out = f.export_network(_random_archive(net, np.random.default_rng(3)), str(tmp_path / "p"), net)
    shipped = params_dir(net)
    for name in os.listdir(shipped):
        if name.endswith(".bin"):
            assert os.path.getsize(os.path.join(out, name)) == os.path.getsize(os.path.join(shipped, name)), name
    o = Oracle(net, out)
    data = synth_cifar(16, 2) if net.startswith("cnv") else synth_mnist(64, 2)
    cls = o.classes(data, 10)
    assert len(set(int(c) for c in cls)) > 1


@pytest.mark.gpu
@pytest.mark.parametrize("net", ["cnvW1A1", "cnvW2A2", "lfcW1A2"])
def test_load_npz_runs_on_the_gpu_like_the_oracle_on_the_exported_directory(net, tmp_path):
    import bnn
    f = _fth()
    bnn.bnn.set_param_dir(os.path.join(GOLDEN, "params"))
    archive = _random_archive(net, np.random.default_rng(11))
    out = f.export_network(archive, str(tmp_path / "p"), net)
    o = Oracle(net, out)
    if net.startswith("cnv"):
        c = bnn.CnvClassifier(net, "cifar10", bnn.RUNTIME_CUDA)
        data = synth_cifar(200, 4)
        f.load_npz(c, archive)
        got = c.classify_records(data)
    else:
        c = bnn.LfcClassifier(net, "mnist", bnn.RUNTIME_CUDA)
        data = synth_mnist(500, 4)
        f.load_npz(c, archive)
        got = c.classify_pixels(data)
    assert np.array_equal(got, o.classes(data, 10))
    assert len(c.classes) == 10


# ---- pinned to the reference exporter's own arithmetic ------------------------------------------------------------------------------
def _ref_cases():
    sys.path.insert(0, os.path.join(GOLDEN))
    import make_golden_exporter as mg
    return mg


def test_bn_folding_equals_reference_functions():
    """bnn.finnthesizer's make_fc_bn / make_fc_bn_qnn / make_conv_bn / quantize against tests/golden/exporter_ref.npz, which
    tests/golden/make_golden_exporter.py produced by executing the reference's own makeFCBNComplex, makeFCBNComplex_QNN,
    makeConvBNComplex and quantize (bnn/src/training/finnthesizer.py:149-330) on the same seeded batch-norm parameters: thresholds
    and (flipped, quantised) weights element for element, float32 and float64 archives, every precision combination the five
    networks' drivers use."""
    mg, f = _ref_cases(), _fth()
    g = np.load(os.path.join(GOLDEN, "exporter_ref.npz"))
    for kind, name, kw in mg.cases():
        w, bias, beta, gamma, mean, invstd = mg.inputs(kind, name, kw)
        if kind == "fc":
            W, T = f.make_fc_bn(w, bias, beta, gamma, mean, invstd, kw["w_int"], kw["w_fract"], use_popcount=kw["popcount"])
        elif kind == "fcq":
            W, T = f.make_fc_bn_qnn(w, bias, beta, gamma, mean, invstd, kw["w_fract"], kw["a_fract"], kw["w_int"], kw["a_int"], kw["tb"], kw["tib"])
        else:
            W, T = f.make_conv_bn(w, bias, beta, gamma, mean, invstd, kw["inter"], kw["w_fract"], kw["a_fract"], 0, kw["w_int"], kw["a_int"], 8,
                                  use_popcount=kw["popcount"], thres_bits=kw["tb"], thres_int_bits=kw["tib"])
        assert np.array_equal(np.asarray(W, np.float64), g[name + "/W"]), name + ": weights"
        assert np.array_equal(np.asarray(T, np.float64).reshape(g[name + "/T"].shape), g[name + "/T"]), name + ": thresholds"
    xs = g["quantize/x"]
    for (i, fr) in ((1, 0), (2, 0), (1, 1), (2, 2)):
        assert np.array_equal(np.asarray(f.quantize(xs, i, fr), np.float64), g["quantize/%d_%d" % (i, fr)]), (i, fr)
