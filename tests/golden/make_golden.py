"""Generates tests/golden/ref_<net>.npz from the UNMODIFIED reference sw runtime (oracle/_ref/python_sw-<net>.so,
built by oracle/build_ref.sh from /root/reference).  Run in the build container only:  python tests/golden/make_golden.py
Each network runs in its own process (the reference keeps its state in process globals)."""
import os
import sys
from multiprocessing import get_context

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

N_CNV, N_LFC, N_FAULT_IMG = 32, 256, 6


def fault_list(net, o, seed=3):
    """Deterministic (target, layer, bit) list: mt19937-free, plain numpy; covers every layer, both memories."""
    rng = np.random.default_rng(seed)
    fl = []
    for l in range(o.nlayers):
        wb, tb = o.weight_layer_bits(l), o.thresh_layer_bits(l)
        for _ in range(6):
            fl.append((1, l, int(rng.integers(0, wb))))
        if tb:
            for _ in range(6):
                fl.append((2, l, int(rng.integers(0, tb))))
    return np.array(fl, np.uint32)


def job(net):
    from oracle_py import Oracle, params_dir, synth_cifar, synth_mnist, mnist_idx_bytes
    from ref_py import Ref
    sys.stdout = open(os.devnull, "w")  # the reference prints a lot
    os.dup2(sys.stdout.fileno(), 1)
    r = Ref(net, params_dir(net))
    o = Oracle(net)  # only for geometry (bit counts) -- results below all come from the reference
    out = {}
    if net.startswith("cnv"):
        deer = open(os.path.join(HERE, "images", "deer.bin"), "rb").read()
        out["deer_scores"] = r.multi(deer, 10, 1)
        out["deer_class_multi"] = r.multi(deer, 10, 0)
        c, s = r.single(deer, 10)
        out["deer_class_single"] = np.array([c])
        out["deer_single_scores"] = s[:10]
        rec = synth_cifar(N_CNV, 0)
        out["synth_scores"] = r.multi(rec.tobytes(), 10, 1)
        out["synth_classes"] = r.multi(rec.tobytes(), 10, 0)
        frec = rec[:N_FAULT_IMG]
    else:
        img3 = open(os.path.join(HERE, "images", "3.image-idx3-ubyte"), "rb").read()
        out["img3_class_multi"] = r.multi(img3, 10, 0)
        c, s = r.single(img3, 10)
        out["img3_class_single"] = np.array([c])
        out["img3_onehot"] = s
        px = synth_mnist(N_LFC, 0)
        out["synth_classes"] = r.multi(mnist_idx_bytes(px), 10, 0)
        # raw 64-bit output words through BlackBoxJam on inputs packed as binarizeAndPack does
        bits = np.zeros((N_LFC, 832), np.uint8)
        bits[:, :784] = px >= 128
        packed = np.packbits(bits, axis=1, bitorder="little").view(np.uint64)
        out["synth_words"] = r.raw(packed, N_LFC, 1)[:, 0]
        frec = px[:64]
    # deterministic faults through the reference's own injector (word_size 1), results after ALL faults applied
    fl = fault_list(net, o)
    out["fault_list"] = fl
    reads = []
    for t, l, b in fl:
        r.inject(0 if t == 1 else 1, int(l), int(b), 1)
    # read back every touched word through FoldedMVMemRead
    for l in range(o.nlayers):
        li = o.layer_info(l)
        reads.append(r.mem_read(2 * l, 0, 0, 0))
        reads.append(r.mem_read(2 * l + 1, li["pe"] - 1, li["tmem"] - 1, 0))
    out["fault_reads"] = np.array(reads, np.uint64)
    if net.startswith("cnv"):
        out["fault_scores"] = r.multi(frec.tobytes(), 10, 1)
    else:
        out["fault_classes"] = r.multi(mnist_idx_bytes(frec), 10, 0)
    # the bool quirk: word_size 8 must behave as 1; flip one more weight bit and read the word
    before = r.mem_read(2, 0, 1, 0)
    r.inject(0, 1, o.layer_info(1)["simd"] * o.layer_info(1)["wbits"] * 1 + 0, 8)
    out["quirk_word"] = np.array([before, r.mem_read(2, 0, 1, 0)], np.uint64)
    np.savez_compressed(os.path.join(HERE, "ref_%s.npz" % net), **out)
    return net


if __name__ == "__main__":
    nets = sys.argv[1:] or ["cnvW1A1", "cnvW1A2", "cnvW2A2", "lfcW1A1", "lfcW1A2"]
    with get_context("spawn").Pool(len(nets)) as p:
        for n in p.imap_unordered(job, nets):
            sys.stderr.write("golden written for %s\n" % n)
