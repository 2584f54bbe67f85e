"""Pins the CPU restatement (oracle/bnn_oracle.c) against vectors produced by the UNMODIFIED reference sw runtime
(tests/golden/ref_<net>.npz, made by tests/golden/make_golden.py) and against the reference's own known-answer tests
(deer.bin -> 4, 3.image-idx3-ubyte -> 3: make-hw.sh:123-129, tests/test_qnn.py:48-133)."""
import os

import numpy as np
import pytest

from oracle_py import GOLDEN, NETS, Oracle, synth_cifar, synth_mnist

CNV = [n for n in NETS if n.startswith("cnv")]
LFC = [n for n in NETS if n.startswith("lfc")]


def gold(net):
    return np.load(os.path.join(GOLDEN, "ref_%s.npz" % net))


def deer():
    return np.frombuffer(open(os.path.join(GOLDEN, "images", "deer.bin"), "rb").read(), np.uint8)


def img3():
    return np.frombuffer(open(os.path.join(GOLDEN, "images", "3.image-idx3-ubyte"), "rb").read(), np.uint8)[16:]


# SURVEY Appendix E (recorded from the reference build during the survey)
APP_E = {
    "cnvW1A1": [228, 229, 251, 250, 412, 253, 230, 262, 222, 239],
    "cnvW1A2": [-26, -44, -32, -8, 264, 2, -12, -32, -46, -36],
    "cnvW2A2": [-22, -27, -16, -10, 239, 1, -9, -17, -15, -27],
}


@pytest.mark.parametrize("net", CNV)
def test_cnv_known_answer_and_scores(net):
    o, g = Oracle(net), gold(net)
    sc = o.cnv_scores(deer())
    assert sc[0, :10].tolist() == APP_E[net]
    assert np.array_equal(sc[:, :10], g["deer_scores"])
    assert o.classes(deer())[0] == 4 == g["deer_class_multi"][0] == g["deer_class_single"][0]
    assert o.L.bo_cnv_class_single(sc[0].ctypes.data, 10) == 4


@pytest.mark.parametrize("net", CNV)
def test_cnv_synthetic_scores(net):
    o, g = Oracle(net), gold(net)
    rec = synth_cifar(g["synth_scores"].shape[0], 0)
    sc = o.cnv_scores(rec)
    assert np.array_equal(sc[:, :10], g["synth_scores"])
    assert np.array_equal(o.classes(rec), g["synth_classes"])


@pytest.mark.parametrize("net", LFC)
def test_lfc_known_answer_and_words(net):
    o, g = Oracle(net), gold(net)
    assert o.classes(img3())[0] == 3 == g["img3_class_multi"][0] == g["img3_class_single"][0]
    onehot = np.zeros(64, np.int32)
    w = o.lfc_words(img3())
    assert o.L.bo_lfc_class_single(int(w[0]), 10, onehot.ctypes.data) == 3
    assert np.array_equal(onehot, g["img3_onehot"])
    px = synth_mnist(g["synth_words"].shape[0], 0)
    mask = np.uint64(0x3FF)
    assert np.array_equal(o.lfc_words(px), g["synth_words"])
    assert np.array_equal(o.classes(px), g["synth_classes"])
    assert (g["synth_words"] & mask != 0).any()


@pytest.mark.parametrize("net", NETS)
def test_faults_match_reference_injector(net):
    """Same (target, layer, bit) list through the reference's inject_fault and through the restatement."""
    o, g = Oracle(net), gold(net)
    for t, l, b in g["fault_list"]:
        o.inject(int(t), int(l), int(b), 1)
    reads = []
    for l in range(o.nlayers):
        li = o.layer_info(l)
        reads.append(o.mem_read(2 * l, 0, 0, 0))
        reads.append(o.mem_read(2 * l + 1, li["pe"] - 1, li["tmem"] - 1, 0))
    assert np.array_equal(np.array(reads, np.uint64), g["fault_reads"])
    if net.startswith("cnv"):
        n = g["fault_scores"].shape[0]
        assert np.array_equal(o.cnv_scores(synth_cifar(32, 0)[:n])[:, :10], g["fault_scores"])
    else:
        n = g["fault_classes"].shape[0]
        assert np.array_equal(o.classes(synth_mnist(256, 0)[:n]), g["fault_classes"])
    # word_size bool quirk (faults.h:42 vs :55): word_size 8 reaches the injector as 1
    before = o.mem_read(2, 0, 1, 0)
    assert before == g["quirk_word"][0]
    li = o.layer_info(1)
    o.inject(1, 1, li["simd"] * li["wbits"], 1)
    assert o.mem_read(2, 0, 1, 0) == g["quirk_word"][1]
    assert bin(int(g["quirk_word"][0] ^ g["quirk_word"][1])).count("1") == 1


def test_short_weight_files_load_zeros():
    """streetview/cnvW1A1 layer 8 and chars_merged/lfcW1A1 layer 3 ship fewer words than WMEM."""
    from oracle_py import params_dir
    o = Oracle("cnvW1A1", params_dir("cnvW1A1", "streetview"))
    li = o.layer_info(8)
    assert o.mem_read(16, 0, li["wmem"] - 1, 0) == 0
    o = Oracle("lfcW1A1", params_dir("lfcW1A1", "chars_merged"))
    assert o.mem_read(6, 0, o.layer_info(3)["wmem"] - 1, 0) == 0


def tap_inputs(net):
    """The inputs of tests/golden/make_golden_taps.py (kept in step with it by test_oracle_taps_vs_reference_templates)."""
    n = 8
    if net.startswith("cnv"):
        rec = synth_cifar(n, 77)
        rec[6, 1:] = ((np.arange(3072, dtype=np.int64) * 7) % 256).astype(np.uint8)
        rec[7] = deer()[:3073]
        return rec
    px = synth_mnist(n, 77)
    px[6] = ((np.arange(784) * 5) % 256).astype(np.uint8)
    px[7] = img3()[:784]
    return px


def tap_gold(net):
    g = np.load(os.path.join(GOLDEN, "taps_%s.npz" % net))
    return g["acc"].astype(np.int32), g["act"]


def wrap_last_cnv_layer(o, acc):
    """CNV layer 8 has no threshold: the reference's accumulator there is the ap_uint<16> score itself (PassThroughActivation<
    ap_uint<16>>, cnvW1A1/hw/top.cpp:232-233), so that layer is compared modulo 2^16."""
    if o.is_cnv:
        acc = acc.copy()
        acc[:, -64:] = acc[:, -64:].astype(np.int16)
    return acc


@pytest.mark.parametrize("net", NETS)
def test_oracle_taps_vs_reference_templates(net):
    """Every layer's accumulators and outputs of the restatement == the reference's own templates replayed layer by layer
    (oracle/tap.cpp: DoCompute's call list with PassThroughActivation for the accumulators), 8 images per network."""
    o = Oracle(net)
    gacc, gact = tap_gold(net)
    data = tap_inputs(net)
    _, acc, act = o.cnv_scores(data, taps=True) if o.is_cnv else o.lfc_words(data, taps=True)
    acc = wrap_last_cnv_layer(o, acc)
    assert acc.shape == gacc.shape and act.shape == gact.shape
    for i in range(data.shape[0]):
        for l, (a, b) in enumerate(zip(o.split_taps(acc[i], "acc"), o.split_taps(gacc[i], "acc"))):
            assert np.array_equal(a, b), "accumulators differ from the reference templates: image %d layer %d" % (i, l)
        for l, (a, b) in enumerate(zip(o.split_taps(act[i], "act"), o.split_taps(gact[i], "act"))):
            assert np.array_equal(a, b), "layer outputs differ from the reference templates: image %d layer %d" % (i, l)
