"""ctypes door to oracle/libbnn_oracle.so (the CPU restatement) -- TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
GOLDEN = os.path.join(ROOT, "tests", "golden")
PARAMS = {
    "cnvW1A1": "cifar10/cnvW1A1", "cnvW1A2": "cifar10/cnvW1A2", "cnvW2A2": "cifar10/cnvW2A2",
    "lfcW1A1": "mnist/lfcW1A1", "lfcW1A2": "mnist/lfcW1A2",
}
NETS = list(PARAMS)


def params_dir(net, dataset=None):
    rel = PARAMS[net] if dataset is None else "%s/%s" % (dataset, net)
    return os.path.join(GOLDEN, "params", rel)


_lib = None


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(ORACLE_DIR, "libbnn_oracle.so")
        src = os.path.join(ORACLE_DIR, "bnn_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "libbnn_oracle.so"])
        L = C.CDLL(so)
        L.bo_open.restype = C.c_void_p
        L.bo_open.argtypes = [C.c_char_p]
        L.bo_close.argtypes = [C.c_void_p]
        L.bo_load.argtypes = [C.c_void_p, C.c_char_p]
        L.bo_num_layers.argtypes = [C.c_void_p]
        L.bo_layer_info.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.bo_mem_set.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint64]
        L.bo_mem_read.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_uint]
        L.bo_mem_read.restype = C.c_uint64
        L.bo_inject.argtypes = [C.c_void_p, C.c_int, C.c_uint, C.c_uint, C.c_uint]
        L.bo_tap_acc_size.argtypes = [C.c_void_p]
        L.bo_tap_acc_size.restype = C.c_size_t
        L.bo_tap_act_size.argtypes = [C.c_void_p]
        L.bo_tap_act_size.restype = C.c_size_t
        L.bo_cifar_quantise.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.bo_cnv_compute.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.bo_lfc_compute.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.bo_cnv_class_multi.argtypes = [C.c_void_p, C.c_int]
        L.bo_cnv_class_single.argtypes = [C.c_void_p, C.c_int]
        L.bo_lfc_class_multi.argtypes = [C.c_uint64, C.c_int]
        L.bo_lfc_class_single.argtypes = [C.c_uint64, C.c_int, C.c_void_p]
        L.bo_weight_layer_bits.argtypes = [C.c_void_p, C.c_int]
        L.bo_weight_layer_bits.restype = C.c_uint
        L.bo_thresh_layer_bits.argtypes = [C.c_void_p, C.c_int]
        L.bo_thresh_layer_bits.restype = C.c_uint
        L.bo_classify_with_faults.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Oracle:
    """One network instance of the CPU restatement."""

    def __init__(self, net, pdir=None):
        self.L = lib()
        self.net = net
        self.h = self.L.bo_open(net.encode())
        assert self.h, net
        self.is_cnv = net.startswith("cnv")
        self.load(pdir or params_dir(net))

    def load(self, pdir):
        assert self.L.bo_load(self.h, pdir.encode()) == 0, pdir

    def __del__(self):
        try:
            self.L.bo_close(self.h)
        except Exception:
            pass

    @property
    def nlayers(self):
        return self.L.bo_num_layers(self.h)

    def layer_info(self, i):
        a = np.zeros(12, np.int32)
        self.L.bo_layer_info(self.h, i, _p(a))
        return dict(zip(["mw", "mh", "ofm_dim", "pool", "out_mode", "in_mode", "pe", "simd", "wmem", "tmem", "nthr", "wbits"], a.tolist()))

    def quantise(self, records):
        records = np.ascontiguousarray(records, np.uint8).reshape(-1, 3073)
        q = np.zeros((records.shape[0], 3072), np.int8)
        self.L.bo_cifar_quantise(_p(records), records.shape[0], _p(q))
        return q

    def cnv_scores(self, records, taps=False):
        """records: n x 3073 uint8.  Returns int16 scores n x 64 (and taps)."""
        q = self.quantise(records)
        n = q.shape[0]
        sc = np.zeros((n, 64), np.int16)
        acc = act = None
        if taps:
            acc = np.zeros((n, self.L.bo_tap_acc_size(self.h)), np.int32)
            act = np.zeros((n, self.L.bo_tap_act_size(self.h)), np.int8)
        assert self.L.bo_cnv_compute(self.h, _p(q), n, _p(sc), _p(acc), _p(act)) == 0
        return (sc, acc, act) if taps else sc

    def lfc_words(self, pixels, taps=False):
        pixels = np.ascontiguousarray(pixels, np.uint8).reshape(-1, 784)
        n = pixels.shape[0]
        w = np.zeros(n, np.uint64)
        acc = act = None
        if taps:
            acc = np.zeros((n, self.L.bo_tap_acc_size(self.h)), np.int32)
            act = np.zeros((n, self.L.bo_tap_act_size(self.h)), np.int8)
        assert self.L.bo_lfc_compute(self.h, _p(pixels), n, _p(w), _p(acc), _p(act)) == 0
        return (w, acc, act) if taps else w

    def classes(self, data, ncls=10):
        if self.is_cnv:
            sc = self.cnv_scores(data)
            return np.array([self.L.bo_cnv_class_multi(_p(sc[i]), ncls) for i in range(sc.shape[0])], np.int32)
        w = self.lfc_words(data)
        return np.array([self.L.bo_lfc_class_multi(int(x), ncls) for x in w], np.int32)

    def split_taps(self, flat, kind):
        """Split one image's flat tap vector into per-layer arrays [rows, mh]."""
        out, o = [], 0
        for i in range(self.nlayers):
            li = self.layer_info(i)
            d = li["ofm_dim"] // 2 if (kind == "act" and li["pool"]) else li["ofm_dim"]
            sz = d * d * li["mh"]
            out.append(flat[o:o + sz].reshape(d * d, li["mh"]))
            o += sz
        return out

    def mem_read(self, tl, mem, ind, th):
        return self.L.bo_mem_read(self.h, tl, mem, ind, th)

    def mem_set(self, tl, mem, ind, th, val):
        self.L.bo_mem_set(self.h, tl, mem, ind, th, val & 0xFFFFFFFFFFFFFFFF)

    def inject(self, target, layer, bit, word_size=1):
        """target: 1 weights, 2 thresholds."""
        self.L.bo_inject(self.h, target, layer, bit, word_size)

    def weight_layer_bits(self, l):
        return self.L.bo_weight_layer_bits(self.h, l)

    def thresh_layer_bits(self, l):
        return self.L.bo_thresh_layer_bits(self.h, l)

    def classify_with_faults(self, data, faults, ncls=10):
        faults = np.ascontiguousarray(faults, np.uint32).reshape(-1, 5)
        if self.is_cnv:
            inp = self.quantise(data)
        else:
            inp = np.ascontiguousarray(data, np.uint8).reshape(-1, 784)
        n = inp.shape[0]
        cls = np.zeros(n, np.int32)
        self.L.bo_classify_with_faults(self.h, _p(inp), n, ncls, _p(faults), faults.shape[0], _p(cls))
        return cls


def synth_cifar(n, seed):
    """label byte 0 + 3072 iid uniform bytes; images 0/1 = all-0 / all-255."""
    rng = np.random.default_rng(seed)
    rec = np.zeros((n, 3073), np.uint8)
    rec[:, 1:] = rng.integers(0, 256, (n, 3072), dtype=np.uint8)
    if n > 0:
        rec[0, 1:] = 0
    if n > 1:
        rec[1, 1:] = 255
    return rec


def synth_mnist(n, seed):
    rng = np.random.default_rng(seed)
    px = rng.integers(0, 256, (n, 784), dtype=np.uint8)
    if n > 0:
        px[0] = 0
    if n > 1:
        px[1] = 255
    return px


def mnist_idx_bytes(px):
    px = np.ascontiguousarray(px, np.uint8).reshape(-1, 784)
    hdr = np.array([2051, px.shape[0], 28, 28], ">u4").tobytes()
    return hdr + px.tobytes()
