"""Generates tests/golden/taps_<net>.npz: every layer's accumulators and outputs for 8 images per network, produced by the
reference's OWN templates (oracle/_ref/tap-<net>.so = oracle/tap.cpp over the unmodified <net>/hw/top.cpp and finn-hlslib,
parameters loaded by the reference's own load_parameters).  Run in the build container only:
    python tests/golden/make_golden_taps.py
Each network runs in its own process (the reference keeps its state in process globals)."""
import ctypes as C
import os
import sys
from multiprocessing import get_context

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))

N_IMG = 8


def tap_inputs(net):
    """The 8 raw inputs: all-0, all-255 (synth images 0/1), four random images, a ramp, and the reference's own test image."""
    from oracle_py import synth_cifar, synth_mnist
    if net.startswith("cnv"):
        rec = synth_cifar(N_IMG, 77)
        ramp = (np.arange(3072, dtype=np.int64) * 7) % 256
        rec[6, 1:] = ramp.astype(np.uint8)
        rec[7] = np.frombuffer(open(os.path.join(HERE, "images", "deer.bin"), "rb").read(), np.uint8)[:3073]
        return rec
    px = synth_mnist(N_IMG, 77)
    px[6] = ((np.arange(784) * 5) % 256).astype(np.uint8)
    px[7] = np.frombuffer(open(os.path.join(HERE, "images", "3.image-idx3-ubyte"), "rb").read(), np.uint8)[16:16 + 784]
    return px


def pack_inputs(net, raw):
    """Raw records / pixels -> the words BlackBoxJam reads (quantiseAndPack<8,1> / binarizeAndPack), via the pinned oracle."""
    from oracle_py import Oracle
    if net.startswith("cnv"):
        return np.ascontiguousarray(Oracle(net).quantise(raw)).view(np.uint64).reshape(raw.shape[0], 384)
    bits = np.zeros((raw.shape[0], 832), np.uint8)
    bits[:, :784] = raw >= 128
    return np.ascontiguousarray(np.packbits(bits, axis=1, bitorder="little")).view(np.uint64).reshape(raw.shape[0], 13)


def run_tap(net, pdir, packed):
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "tap-%s.so" % net))
    L.load_parameters.argtypes = [C.c_char_p]
    L.ref_tap_acc_size.restype = C.c_size_t
    L.ref_tap_act_size.restype = C.c_size_t
    L.ref_tap.argtypes = [C.c_void_p, C.c_uint, C.c_void_p, C.c_void_p]
    L.load_parameters(pdir.encode())
    n = packed.shape[0]
    acc = np.zeros((n, L.ref_tap_acc_size()), np.int32)
    act = np.zeros((n, L.ref_tap_act_size()), np.int8)
    assert L.ref_tap(packed.ctypes.data_as(C.c_void_p), n, acc.ctypes.data_as(C.c_void_p), act.ctypes.data_as(C.c_void_p)) == 0
    return acc, act


def job(net):
    from oracle_py import params_dir
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)  # the reference prints while loading
    raw = tap_inputs(net)
    acc, act = run_tap(net, params_dir(net), pack_inputs(net, raw))
    assert np.abs(acc).max() < 32768
    np.savez_compressed(os.path.join(HERE, "taps_%s.npz" % net), acc=acc.astype(np.int16), act=act)
    return net


if __name__ == "__main__":
    nets = sys.argv[1:] or ["cnvW1A1", "cnvW1A2", "cnvW2A2", "lfcW1A1", "lfcW1A2"]
    with get_context("spawn").Pool(len(nets)) as p:
        for n in p.imap_unordered(job, nets):
            sys.stderr.write("golden taps written for %s\n" % n)
