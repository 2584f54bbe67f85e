"""Trained-network exporter for the CUDA runtime: the counterpart of BNN-PYNQ's bnn/src/training/finnthesizer.py and of its
per-network drivers (cifar10-gen-weights-W1A1/W1A2/W2A2.py, mnist-gen-weights-W1A1/W1A2.py).

Input: the .npz archive the reference's training scripts save (arr_0, arr_1, ... = per layer: weights, bias, beta, gamma, mean,
invstd; the last CNV layer has weights only).  Output: a parameter directory `<L>-<PE>-weights.bin` / `<L>-<PE>-thres.bin` /
`classes.txt`, byte-identical to what the reference exporter writes, i.e. exactly what `load_parameters`
reads — or, with `load_npz`, straight into a classifier without keeping the directory.

Written from the behaviour of the reference, not from its code: the reference is Python 2, loops over every synapse and formats
words through strings and the `bitstring` package; here every step is a numpy array operation.
  thresholds      makeFCBNComplex :167-205, makeFCBNComplex_QNN :211-256, makeConvBNComplex :260-327
  reader          BNNWeightReader :332-442 (archive order, channel interleave of the first FC layer after the convolutions)
  padding/folding BNNProcElemMem.__padMatrix :523-551, __updatePEMapping :553-588 (row r -> PE r % PE, tile nf*SF + sf)
  word formats    binArrayToString :461-463 (first column = LSB), ArrayToString :468-482 (two's complement fields), __wmem2bin /
                  __tmem2bin :616-690 (little-endian 64-bit words, int16 saturation, pad threshold 65535 -> 32767)
  drivers         convertFCNetwork :39-104 (first layer padded to 64 columns, last to 64 rows), cifar10-gen-weights-W1A1.py:44-186
The config.h / HLS initialiser outputs of the reference describe the FPGA build and are not produced.
"""
import os
import tempfile

import numpy as np

THRES_BITS = 16                      # BNNProcElemMem's default numThresBits; the drivers never override it


def pad_to(value, multiple):
    return value if value % multiple == 0 else value + multiple - value % multiple


def quantize(x, integer, fract):
    """1 bit: 1 where x >= 0 else 0 (binarize, :160-161).  Otherwise round-half-up to `fract` fractional bits, no clipping (:152-157)."""
    x = np.asarray(x, dtype=np.float64)
    if integer + fract == 1:
        return (x >= 0).astype(np.int64)
    n = float(2 ** fract)
    return np.floor(x * n + 0.5) / n


def _activation_steps(a_int, a_fract):
    """Quantisation steps of the hardtanh activation: for 2 bits [-0.5, +0.5] (:223, :275)."""
    return np.linspace(-1, 1, num=2 ** (a_int + a_fract - 1), endpoint=False) + 1.0 / (2 ** (a_fract + 1))


def _trunc(x):
    """Python's int(): toward zero."""
    return np.trunc(x).astype(np.int64)


def _signed_weights(weights_out_major, flip, w_int, w_fract):
    """Quantised weights with the rows of negative-slope neurons negated BEFORE quantisation (so binarize(-0.0) = 1 as in the reference)."""
    w = np.asarray(weights_out_major, dtype=np.float64)
    w = np.where(flip[:, None], -w, w)
    q = quantize(w, w_int, w_fract)
    return q.astype(np.int64) if w_fract == 0 else q


def make_fc_bn(weights, bias, beta, gamma, mean, invstd, w_int=1, w_fract=0, use_popcount=True):
    """Single-threshold fully connected layer.  weights [ins, outs] -> (W [outs, ins], T [outs]).
    threshold on the signed sum: mean - bias - beta / (gamma*invstd), negated together with the weights where gamma*invstd < 0;
    for the XNOR-popcount datapath it becomes the count of ones, int((ins + thres) / 2)."""
    weights = np.asarray(weights, dtype=np.float64)
    ins = weights.shape[0]
    slope = np.asarray(gamma) * np.asarray(invstd)                              # in the archive's own precision, as the reference does
    thres = (np.asarray(mean) - np.asarray(bias) - np.asarray(beta) / slope).astype(np.float64)
    flip = slope < 0
    thres = np.where(flip, -thres, thres)
    T = _trunc((ins + thres) / 2) if use_popcount else _trunc(thres)
    return _signed_weights(weights.T, flip, w_int, w_fract), T


def make_fc_bn_qnn(weights, bias, beta, gamma, mean, invstd, w_fract, a_fract, w_int, a_int, thres_bits=THRES_BITS, thres_int_bits=None):
    """Multi-threshold fully connected layer: one threshold per activation step, floor (ceil of the negation where flipped)."""
    weights = np.asarray(weights, dtype=np.float64)
    step = _activation_steps(a_int, a_fract)
    factor = 1 if thres_int_bits is None else 2 ** (thres_bits - thres_int_bits)
    slope = (np.asarray(gamma) * np.asarray(invstd)).astype(np.float64)[:, None]  # product and difference in the archive's precision,
    base = (np.asarray(mean) - np.asarray(bias)).astype(np.float64)[:, None]      # the rest in double (numpy promotion in the reference)
    thres = base + (step[None, :] - np.asarray(beta, np.float64)[:, None]) / slope
    flip = slope[:, 0] < 0
    T = np.where(flip[:, None], np.ceil(-factor * thres), np.floor(factor * thres)).astype(np.int64)
    return _signed_weights(weights.T, flip, w_int, w_fract), T


def make_conv_bn(weights, bias, beta, gamma, mean, invstd, interleave_channels, w_fract, a_fract, i_fract, w_int, a_int, i_int,
                 use_popcount=True, thres_bits=THRES_BITS, thres_int_bits=None):
    """Convolution + batch-norm.  weights [ofm, ifm, k, k] -> (W [ofm, k*k*ifm], T [ofm] or [ofm, steps]).
    Column order (ky, kx, ifm) when interleave_channels (what the sliding-window unit produces), else (ifm, ky, kx)."""
    weights = np.asarray(weights, dtype=np.float64)
    ofm, ifm, k, k2 = weights.shape
    if k != k2:
        raise ValueError("non-square convolution kernels are not supported")
    fanin = ifm * k * k
    factor = 1 if thres_int_bits is None else 2 ** (thres_bits - thres_int_bits)
    slope = np.asarray(gamma) * np.asarray(invstd)                              # archive precision (see make_fc_bn)
    flip = slope < 0
    base = np.asarray(mean) - np.asarray(bias)
    if a_int + a_fract == 1:
        thres = (base - np.asarray(beta) / slope).astype(np.float64)
        thres = np.where(flip, np.ceil(-factor * thres), np.floor(factor * thres))
        T = _trunc((fanin + thres) / 2) if use_popcount else thres
    else:
        step = _activation_steps(a_int, a_fract)
        thres = base.astype(np.float64)[:, None] + (step[None, :] - np.asarray(beta, np.float64)[:, None]) / slope.astype(np.float64)[:, None]
        T = np.where(flip[:, None], np.ceil(-factor * thres), np.floor(factor * thres))
    cols = weights.transpose(0, 2, 3, 1) if interleave_channels else weights
    return _signed_weights(cols.reshape(ofm, fanin), flip, w_int, w_fract), T


class WeightReader:
    """Walks the arrays of a training archive in order (arr_0, arr_1, ...)."""

    def __init__(self, param_file, interleave_channels):
        self.params = param_file if isinstance(param_file, dict) else np.load(param_file)
        self.index = 0
        self.interleave_channels = interleave_channels
        self.num_interleave_channels = 0

    def _next(self):
        a = self.params["arr_%d" % self.index]
        self.index += 1
        return a

    def read_weights_raw(self):
        return self._next()

    def read_batch_norm_raw(self):
        return tuple(self._next() for _ in range(5))        # bias, beta, gamma, mean, invstd

    def _fc(self, weights, bn, w_fract, a_fract, i_fract, w_int, a_int, i_int, thres_bits, thres_int_bits):
        if w_fract + w_int == 1 and a_fract + a_int == 1 and i_fract + i_int == 1:
            W, T = make_fc_bn(weights, *bn, w_int, w_fract, use_popcount=True)
        elif a_fract + a_int == 1:
            W, T = make_fc_bn(weights, *bn, w_int, w_fract, use_popcount=False)
        else:
            W, T = make_fc_bn_qnn(weights, *bn, w_fract, a_fract, w_int, a_int, thres_bits, thres_int_bits)
        if self.interleave_channels and self.num_interleave_channels:
            # first FC layer after the convolutions: columns from (channel, pixel) to (pixel, channel) order
            c = self.num_interleave_channels
            W = W.reshape(W.shape[0], c, W.shape[1] // c).transpose(0, 2, 1).reshape(W.shape[0], -1)
            self.num_interleave_channels = 0
        return W, T

    def read_fc_bn(self, w_fract, a_fract, i_fract, w_int, a_int, i_int, thres_bits=THRES_BITS, thres_int_bits=None):
        weights = self.read_weights_raw()
        return self._fc(weights, self.read_batch_norm_raw(), w_fract, a_fract, i_fract, w_int, a_int, i_int, thres_bits, thres_int_bits)

    def read_fc_no_thresholds(self, w_fract, a_fract, i_fract, w_int, a_int, i_int, thres_bits=THRES_BITS, thres_int_bits=None):
        weights = self.read_weights_raw()
        n = weights.shape[1]
        bn = (np.zeros(n), np.zeros(n), np.ones(n), np.ones(n), np.ones(n))
        return self._fc(weights, bn, w_fract, a_fract, i_fract, w_int, a_int, i_int, thres_bits, thres_int_bits)

    def read_conv_bn(self, w_fract, a_fract, i_fract, w_int, a_int, i_int, use_popcount=True, thres_bits=THRES_BITS, thres_int_bits=None):
        weights = self.read_weights_raw()
        bn = self.read_batch_norm_raw()
        self.num_interleave_channels = weights.shape[0]
        return make_conv_bn(weights, *bn, self.interleave_channels, w_fract, a_fract, i_fract, w_int, a_int, i_int,
                            use_popcount=use_popcount, thres_bits=thres_bits, thres_int_bits=thres_int_bits)


class ProcElemMem:
    """Per-PE weight and threshold memories of one engine, and their on-disk form.
    Row r of a padded matrix goes to PE r % PE; within a PE rows are stored in order, each as SF = cols / SIMD words whose field s
    (LSB first) is column sf*SIMD + s."""

    def __init__(self, num_pe, num_simd, weight_mem_depth, thres_mem_depth, w_int, a_int, i_int, w_fract=0, a_fract=0, i_fract=0,
                 thres_bits=THRES_BITS, thres_int_bits=None):
        self.num_pe, self.num_simd = int(num_pe), int(num_simd)
        self.w_prec, self.a_prec, self.i_prec = w_int + w_fract, a_int + a_fract, i_int + i_fract
        if self.num_simd * self.w_prec > 64:
            raise ValueError("SIMD x weight bits must fit the 64-bit memory word")
        if self.a_prec > 2:
            raise ValueError("activations wider than 2 bits are not supported by the runtime")
        self.thres_bits, self.thres_int_bits = thres_bits, thres_int_bits
        self.num_thresholds = 2 ** (self.a_prec - 1)
        pad_word = (1 << self.num_simd) - 1 if self.w_prec == 1 else 0           # unused entries: all-ones bits / zero weights
        self.weight_mem = np.full((self.num_pe, int(weight_mem_depth)), pad_word, dtype=np.uint64)
        self.thres_mem = np.zeros((self.num_pe, int(thres_mem_depth), self.num_thresholds), dtype=np.float64)
        self.weight_head = 0
        self.thres_head = 0
        self.accu_offset = 0
        self.layer_sizes, self.neuron_pad, self.synapse_pad = [], [], []

    def _pad(self, W, T, pad_w, pad_h):
        n, s = W.shape
        pad_n, pad_s = pad_h - n, pad_w - s
        if pad_n < 0 or pad_s < 0:
            raise ValueError("padded shape smaller than the matrix")
        Wp = np.pad(W, ((0, pad_n), (0, pad_s)), "constant", constant_values=1 if self.w_prec == 1 else 0)   # input pad bits are 0
        T = np.asarray(T, dtype=np.float64).reshape(n, -1)
        if T.shape[1] != self.num_thresholds:
            raise ValueError("expected %d threshold(s) per neuron, got %d" % (self.num_thresholds, T.shape[1]))
        Tp = np.pad(T, ((0, pad_n), (0, 0)), "constant", constant_values=2 ** self.thres_bits - 1)
        self.neuron_pad.append(pad_n)
        self.synapse_pad.append(pad_s)
        # for datapaths where a pad column is not neutral (+1 weight against a 0 input read as -1 ... or as 0), the reference moves
        # the pad contribution into the thresholds; XNOR-popcount and multi-bit weights need nothing
        transparent = (self.w_prec == 1 and self.a_prec == 1 and self.i_prec == 1) or self.w_prec >= 2
        self.accu_offset = 0 if transparent else pad_s
        return Wp, Tp

    def add_matrix(self, W, T, pad_w, pad_h):
        if self.thres_int_bits is not None:                                     # fixed-point thresholds: keep the representable part
            sh = 2.0 ** (self.thres_bits - self.thres_int_bits)
            T = np.trunc(np.asarray(T, np.float64) * (sh if self.a_prec == 1 else 1)) / sh
        Wp, Tp = self._pad(np.asarray(W), T, pad_w, pad_h)
        n, s = Wp.shape
        if n % self.num_pe or s % self.num_simd:
            raise ValueError("padded matrix must be a multiple of PE x SIMD")
        sf, nf = s // self.num_simd, n // self.num_pe
        fields = np.trunc(Wp).astype(np.int64).reshape(n, sf, self.num_simd)
        if self.w_prec == 1:
            if ((fields != 0) & (fields != 1)).any():
                raise ValueError("non-binary values in 1-bit weight data")
        else:
            fields = np.where(fields < 0, fields + (1 << self.w_prec), fields)
        shifts = (np.arange(self.num_simd, dtype=np.uint64) * np.uint64(self.w_prec))[None, None, :]
        words = np.bitwise_or.reduce(fields.astype(np.uint64) << shifts, axis=2)                       # [n, sf]
        if self.weight_head + nf * sf > self.weight_mem.shape[1] or self.thres_head + nf > self.thres_mem.shape[1]:
            raise ValueError("matrix does not fit the PE memories")
        # rows r = f*PE + p  ->  PE p, entries [head + f*sf, head + (f+1)*sf)
        self.weight_mem[:, self.weight_head:self.weight_head + nf * sf] = words.reshape(nf, self.num_pe, sf).transpose(1, 0, 2).reshape(self.num_pe, nf * sf)
        self.thres_mem[:, self.thres_head:self.thres_head + nf, :] = (Tp - self.accu_offset).reshape(nf, self.num_pe, -1).transpose(1, 0, 2)
        self.weight_head += nf * sf
        self.thres_head += nf
        self.layer_sizes.append((n, s))

    def threshold_words(self):
        """int64 [PE][entries][thresholds] as written: saturated to the signed threshold width, then truncated toward zero."""
        if self.thres_int_bits is None:
            lim = 2 ** (self.thres_bits - 1)
            return np.trunc(np.clip(self.thres_mem, -lim, lim - 1)).astype(np.int64)
        return np.trunc(self.thres_mem * 2.0 ** (self.thres_bits - self.thres_int_bits)).astype(np.int64)

    def write_bin_files(self, target_dir, prefix="", use_thresholds=True):
        t = self.threshold_words()
        for pe in range(self.num_pe):
            self.weight_mem[pe].astype("<u8").tofile(os.path.join(target_dir, "%s-%d-weights.bin" % (prefix, pe)))
            if use_thresholds:
                t[pe].astype("<i8").tofile(os.path.join(target_dir, "%s-%d-thres.bin" % (prefix, pe)))


# Folding and precisions of the five networks (peCounts / simdCounts / *Precisions_* of the reference's drivers; the same numbers as
# bnn/src/network/<net>/hw/config.h).  Fractional bits are 0 everywhere except the 8-bit CNV input (1.7).
_CNV_BASE = dict(kind="cnv", pe=[16, 32, 16, 16, 4, 1, 1, 1, 4], simd=[3, 32, 32, 32, 32, 32, 4, 8, 1], i_fract=[7] + [0] * 8)
NETWORKS = {
    "cnvW1A1": dict(_CNV_BASE, w_int=[1] * 9, a_int=[1] * 9, i_int=[1] * 9),
    "cnvW1A2": dict(_CNV_BASE, w_int=[1] * 9, a_int=[2] * 8 + [1], i_int=[1] + [2] * 8),
    "cnvW2A2": dict(_CNV_BASE, pe=[8, 16, 8, 8, 4, 1, 1, 2, 4], simd=[3, 16, 16, 16, 8, 8, 2, 2, 1], w_int=[2] * 9, a_int=[2] * 8 + [1],
                    i_int=[1] + [2] * 8),
    "lfcW1A1": dict(kind="lfc", simd=[64, 32, 64, 8], pe=[32, 64, 32, 16], i_fract=[0] * 4, w_int=[1] * 4, a_int=[1] * 4, i_int=[1] * 4),
    "lfcW1A2": dict(kind="lfc", simd=[64, 32, 64, 8], pe=[32, 64, 32, 16], i_fract=[0] * 4, w_int=[1] * 4, a_int=[2, 2, 2, 1], i_int=[1, 2, 2, 2]),
}
CIFAR10_CLASSES = ["Airplane", "Automobile", "Bird", "Cat", "Deer", "Dog", "Frog", "Horse", "Ship", "Truck"]


def _write_classes(target_dir, classes):
    with open(os.path.join(target_dir, "classes.txt"), "w") as f:
        f.write("\n".join(str(c) for c in classes))


def convert_fc_network(npz_file, target_dir, simd_counts, pe_counts, w_fract, a_fract, i_fract, w_int, a_int, i_int):
    """Fully connected network (convertFCNetwork :39-104): first layer padded to a multiple of 64 columns, last to 64 rows (the AXI
    word); every layer, the last included, gets threshold files."""
    os.makedirs(target_dir, exist_ok=True)
    r = WeightReader(npz_file, False)
    n_layers = len(simd_counts)
    for l in range(n_layers):
        simd, pe = simd_counts[l], pe_counts[l]
        W, T = r.read_fc_bn(w_fract[l], a_fract[l], i_fract[l], w_int[l], a_int[l], i_int[l])
        padded_h = pad_to(W.shape[0], max(pe, 64) if l == n_layers - 1 else pe)
        padded_w = pad_to(W.shape[1], max(simd, 64) if l == 0 else simd)
        m = ProcElemMem(pe, simd, padded_w * padded_h // (simd * pe), padded_h // pe, w_int[l], a_int[l], i_int[l], w_fract[l], a_fract[l], i_fract[l])
        m.add_matrix(W, T, padded_w, padded_h)
        m.write_bin_files(target_dir, str(l))


def convert_cnv_network(npz_file, target_dir, simd_counts, pe_counts, w_fract, a_fract, i_fract, w_int, a_int, i_int):
    """CNV (cifar10-gen-weights-*.py): six convolutions with channel-interleaved columns — the first with 8-bit inputs, signed-sum
    thresholds in 1/256 units — then three fully connected layers, the last without thresholds and padded to 64 rows."""
    os.makedirs(target_dir, exist_ok=True)
    r = WeightReader(npz_file, True)
    for l in range(9):
        simd, pe = simd_counts[l], pe_counts[l]
        prec = (w_fract[l], a_fract[l], i_fract[l], w_int[l], a_int[l], i_int[l])
        use_thresholds = True
        if l == 0:
            W, T = r.read_conv_bn(*prec, use_popcount=False, thres_bits=24, thres_int_bits=16)
        elif l < 6:
            W, T = r.read_conv_bn(*prec)
        elif l < 8:
            W, T = r.read_fc_bn(*prec)
        else:
            W, T = r.read_fc_no_thresholds(*prec)
            use_thresholds = False
        padded_h = pad_to(W.shape[0], 64 if l == 8 else pe)
        padded_w = pad_to(W.shape[1], simd)
        m = ProcElemMem(pe, simd, padded_w * padded_h // (simd * pe), padded_h // pe, w_int[l], a_int[l], i_int[l], w_fract[l], a_fract[l], i_fract[l])
        m.add_matrix(W, T, padded_w, padded_h)
        m.write_bin_files(target_dir, str(l), use_thresholds)


def export_network(npz_file, target_dir, network, classes=None):
    """npz archive -> parameter directory for `network` (one of NETWORKS), with the folding the runtime's loader expects."""
    cfg = NETWORKS[network]
    n = len(cfg["pe"])
    zeros = [0] * n
    convert = convert_cnv_network if cfg["kind"] == "cnv" else convert_fc_network
    convert(npz_file, target_dir, cfg["simd"], cfg["pe"], zeros, zeros, cfg["i_fract"], cfg["w_int"], cfg["a_int"], cfg["i_int"])
    if classes is None:
        classes = CIFAR10_CLASSES if cfg["kind"] == "cnv" else [str(i) for i in range(10)]
    _write_classes(target_dir, classes)
    return target_dir


def load_npz(classifier, npz_file, classes=None):
    """Export `npz_file` for the classifier's network into a scratch directory and load it: trained archive -> device, no parameter
    directory to keep.  `classifier` is a CnvClassifier / LfcClassifier (or anything with .net and .bnn.load_parameters)."""
    with tempfile.TemporaryDirectory(prefix="bnn_b200_params_") as d:
        export_network(npz_file, d, classifier.net, classes)
        classifier.bnn.load_parameters(d)
    classifier.classes = classifier.bnn.classes
    return classifier
