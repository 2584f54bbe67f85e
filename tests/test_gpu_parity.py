"""GPU parity tests proper: the CUDA path, called through the C ABI (libbnn_b200.so / python_cuda-<net>.so), against the
CPU oracle on the same seeded inputs and against the committed golden vectors of the reference runtime.  Bit-exact."""
import os
import sys

import numpy as np
import pytest

from oracle_py import GOLDEN, NETS, Oracle, mnist_idx_bytes, params_dir, synth_cifar, synth_mnist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))

pytestmark = pytest.mark.gpu
CNV = [n for n in NETS if n.startswith("cnv")]
LFC = [n for n in NETS if n.startswith("lfc")]


def runtime(net, pdir=None):
    from bnn import native
    return native.Runtime(net, pdir or params_dir(net), devices=[0])


def gold(net):
    return np.load(os.path.join(GOLDEN, "ref_%s.npz" % net))


def data_for(net, n, seed):
    return synth_cifar(n, seed) if net.startswith("cnv") else synth_mnist(n, seed)


@pytest.mark.parametrize("net", CNV)
def test_cnv_scores_vs_oracle_and_golden(net):
    rt, o, g = runtime(net), Oracle(net), gold(net)
    deer = np.frombuffer(open(os.path.join(GOLDEN, "images", "deer.bin"), "rb").read(), np.uint8)
    cls, sc = rt.classify_cifar(deer, 10, want_scores=True)
    assert cls[0] == 4 and np.array_equal(sc[:, :10], g["deer_scores"])
    rec = synth_cifar(32, 0)
    cls, sc = rt.classify_cifar(rec, 10, want_scores=True)
    assert np.array_equal(sc[:, :10], g["synth_scores"]) and np.array_equal(cls, g["synth_classes"])
    rec = synth_cifar(600, 11)  # ragged vs the 4-row groups and the chunking
    cls, sc = rt.classify_cifar(rec, 10, want_scores=True)
    assert np.array_equal(sc, o.cnv_scores(rec))
    assert np.array_equal(cls, o.classes(rec))


@pytest.mark.parametrize("net", LFC)
def test_lfc_words_vs_oracle_and_golden(net):
    rt, o, g = runtime(net), Oracle(net), gold(net)
    img3 = np.frombuffer(open(os.path.join(GOLDEN, "images", "3.image-idx3-ubyte"), "rb").read(), np.uint8)[16:]
    assert rt.classify_mnist(img3)[0] == 3
    px = synth_mnist(256, 0)
    cls, words = rt.classify_mnist(px, 10, want_words=True)
    assert np.array_equal(words, g["synth_words"]) and np.array_equal(cls, g["synth_classes"])
    px = synth_mnist(5001, 12)
    cls, words = rt.classify_mnist(px, 10, want_words=True)
    assert np.array_equal(words, o.lfc_words(px)) and np.array_equal(cls, o.classes(px))
    for ncls in (1, 7, 47, 64):  # mask widths of the LFC decode
        assert np.array_equal(rt.classify_mnist(px[:300], ncls), o.classes(px[:300], ncls))


def structured(net, n=6):
    """Inputs with structure instead of iid bytes: constant blocks, ramps, one saturated channel, a checkerboard."""
    if net.startswith("cnv"):
        rec = np.zeros((n, 3073), np.uint8)
        img = rec[:, 1:].reshape(n, 3, 32, 32)
        img[0, :, :16, :] = 255                                         # half black, half white
        img[1] = (np.arange(32, dtype=np.int64)[None, None, :] * 8).astype(np.uint8)      # horizontal ramp
        img[2] = (np.arange(32, dtype=np.int64)[None, :, None] * 8 + 3).astype(np.uint8)  # vertical ramp
        img[3, 0] = 255                                                 # red channel saturated, others zero
        img[4] = ((np.add.outer(np.arange(32), np.arange(32)) % 2) * 255).astype(np.uint8)[None]  # checkerboard
        img[5] = 128                                                    # mid grey: quantises to the value nearest zero
        return rec
    px = np.zeros((n, 784), np.uint8)
    im = px.reshape(n, 28, 28)
    im[0, :14] = 255
    im[1] = (np.arange(28, dtype=np.int64)[None, :] * 9).astype(np.uint8)
    im[2] = (np.arange(28, dtype=np.int64)[:, None] * 9 + 4).astype(np.uint8)
    im[3, 10:18, 10:18] = 255
    im[4] = ((np.add.outer(np.arange(28), np.arange(28)) % 2) * 255).astype(np.uint8)
    im[5] = 128
    return px


def oracle_taps(o, data):
    _, oacc, oact = o.cnv_scores(data, taps=True) if o.is_cnv else o.lfc_words(data, taps=True)
    return oacc, oact


def check_taps(o, acc, act, oacc, oact, what):
    for i in range(acc.shape[0]):
        for l, (a, b) in enumerate(zip(o.split_taps(acc[i], "acc"), o.split_taps(oacc[i], "acc"))):
            if o.is_cnv and l == o.nlayers - 1:   # no threshold: the reference's accumulator IS the 16-bit score (top.cpp:232-233)
                a, b = a.astype(np.int16), b.astype(np.int16)
            assert np.array_equal(a, b), "%s: accumulators differ: image %d layer %d" % (what, i, l)
        for l, (a, b) in enumerate(zip(o.split_taps(act[i], "act"), o.split_taps(oact[i], "act"))):
            assert np.array_equal(a, b), "%s: layer outputs differ: image %d layer %d" % (what, i, l)


@pytest.mark.parametrize("core", [2, 1, 0])
@pytest.mark.parametrize("net", NETS)
def test_every_layer_accumulators_and_outputs(net, core):
    """north_star: every layer's accumulators and threshold outputs, bit-exact, ON EVERY CORE.  Cores 1 / 2 run the TAP
    instantiations of the tensor-core kernels that ship (the fused layer-0/1 kernel, conv_f4 / gemm_f4 or conv_tc / gemm_tc) with
    the production activation formats; the oracle's taps are pinned to the reference's own templates (tests/golden/taps_<net>.npz)."""
    rt, o = runtime(net), Oracle(net)
    rt.set_core(core)
    data = np.concatenate([data_for(net, 17, 21), structured(net)])  # 17: more than one 16-image tap step; + structured inputs
    acc, act = rt.tap(data)
    oacc, oact = oracle_taps(o, data)
    check_taps(o, acc, act, oacc, oact, "core %d" % core)


@pytest.mark.parametrize("core", [2, 0])
@pytest.mark.parametrize("net", NETS)
def test_every_layer_vs_reference_template_taps(net, core):
    """The same taps against the committed per-layer tensors the reference's own templates produced (oracle/tap.cpp)."""
    from test_oracle_golden import tap_gold, tap_inputs
    rt, o = runtime(net), Oracle(net)
    rt.set_core(core)
    gacc, gact = tap_gold(net)
    acc, act = rt.tap(tap_inputs(net))
    check_taps(o, acc, act, gacc, gact, "core %d vs reference templates" % core)


@pytest.mark.parametrize("core", [2, 1, 0])
@pytest.mark.parametrize("net", NETS)
def test_structured_inputs(net, core):
    rt, o = runtime(net), Oracle(net)
    rt.set_core(core)
    d = structured(net)
    if net.startswith("cnv"):
        cls, sc = rt.classify_cifar(d, 10, want_scores=True)
        assert np.array_equal(sc, o.cnv_scores(d))
    else:
        cls, words = rt.classify_mnist(d, 10, want_words=True)
        assert np.array_equal(words, o.lfc_words(d))
    assert np.array_equal(cls, o.classes(d))


@pytest.mark.parametrize("net", NETS)
def test_prepared_inputs_equal_raw_inputs(net):
    """bnnb_classify_packed: inputs in BlackBoxJam's own format (LFC 104-byte binarizeAndPack words, CNV int8 HWC) give the results
    of the raw-input call, through the chunk pipeline (ragged size) and with an empty batch."""
    from bnn import native
    rt, o = runtime(net), Oracle(net)
    d = data_for(net, 9001, 17)
    if net.startswith("cnv"):
        packed = o.quantise(d)                       # the pinned oracle's quantiseAndPack<8,1>
        cls, raw = rt.classify_packed(packed, 10, want_raw=True)
        wcls, wraw = rt.classify_cifar(d, 10, want_scores=True)
        assert np.array_equal(raw[:512], o.cnv_scores(d[:512]))
    else:
        packed = native.pack_mnist(d)
        assert packed.shape == (9001, 104)
        cls, raw = rt.classify_packed(packed, 10, want_raw=True)
        wcls, wraw = rt.classify_mnist(d, 10, want_words=True)
        assert np.array_equal(raw, o.lfc_words(d))
    assert np.array_equal(raw, wraw) and np.array_equal(cls, wcls)
    assert rt.classify_packed(packed[:0], 10).shape == (0,)


@pytest.mark.parametrize("net", ["cnvW1A1", "lfcW1A1"])
def test_small_calls_take_little_memory_and_release_frees_it(net):
    """A one-image call sets up ONE device with buffers for one small chunk; release() (what deinit() calls) frees every device
    allocation and the next call rebuilds the state with the same parameters and faults (rawhls-offload.cpp:54-74)."""
    from bnn import native
    rt = native.Runtime(net, params_dir(net))           # every visible device is a candidate
    o = Oracle(net)
    d = data_for(net, 3, 8)
    assert rt.device_bytes(0) == 0                        # load_parameters allocates nothing by itself
    assert np.array_equal(rt.classify(d[:1]), o.classes(d[:1]))
    assert 0 < rt.device_bytes(0) < 64 << 20, rt.device_bytes(0)
    for slot in range(1, rt.L.bnnb_num_devices(rt.h)):
        assert rt.device_bytes(slot) == 0                 # a small batch stays on one device
    rt.inject_fault(1, 1, 5, 1)
    o.inject(1, 1, 5, 1)
    rt.release()
    assert rt.device_bytes(0) == 0
    assert np.array_equal(rt.classify(d), o.classes(d))   # re-initialised on demand, fault kept
    assert rt.mem_read(2, 0, 0, 0) == o.mem_read(2, 0, 0, 0)


@pytest.mark.parametrize("net", NETS)
def test_edge_cases(net):
    rt, o = runtime(net), Oracle(net)
    assert rt.classify(data_for(net, 0, 0)).shape == (0,)          # empty batch
    for n in (1, 2, 3, 5):                                         # ragged tails
        d = data_for(net, n, 30 + n)
        assert np.array_equal(rt.classify(d), o.classes(d))


@pytest.mark.parametrize("net", NETS)
def test_faults_vs_reference_golden_and_oracle(net):
    rt, o, g = runtime(net), Oracle(net), gold(net)
    for t, l, b in g["fault_list"]:
        rt.inject_fault(int(t), int(l), int(b), 1)
        o.inject(int(t), int(l), int(b), 1)
    reads = []
    for l in range(o.nlayers):
        li = o.layer_info(l)
        reads += [rt.mem_read(2 * l, 0, 0, 0), rt.mem_read(2 * l + 1, li["pe"] - 1, li["tmem"] - 1, 0)]
    assert np.array_equal(np.array(reads, np.uint64), g["fault_reads"])
    if net.startswith("cnv"):
        n = g["fault_scores"].shape[0]
        _, sc = rt.classify_cifar(synth_cifar(32, 0)[:n], 10, want_scores=True)
        assert np.array_equal(sc[:, :10], g["fault_scores"])
    else:
        n = g["fault_classes"].shape[0]
        assert np.array_equal(rt.classify_mnist(synth_mnist(256, 0)[:n]), g["fault_classes"])
    d = data_for(net, 200, 5)
    assert np.array_equal(rt.classify(d), o.classes(d))             # still equal to the (equally faulted) oracle
    rt.load_parameters(params_dir(net))                             # a reload discards the faults
    assert np.array_equal(rt.classify(d), Oracle(net).classes(d))


@pytest.mark.parametrize("net", NETS)
def test_fault_epochs_match_one_image_at_a_time(net):
    """Scheduled faults: time i applies before image i, cumulatively (faults.h:138-146)."""
    rt, o = runtime(net), Oracle(net)
    n = 96
    d = data_for(net, n, 6)
    fl = rt.schedule_faults(n, 40, word_size=1, target=-1, seed=3)
    fl[:3, 0] = [0, 0, n - 1]  # faults before the first image and at the last one
    assert fl.shape[0] > 30 and set(fl[:, 1]) <= {1, 2}
    got = rt.classify_with_fault_list(d, fl)
    want = o.classify_with_faults(d, fl[np.argsort(fl[:, 0], kind="stable")])
    assert np.array_equal(got, want)
    assert np.array_equal(rt.classify(d), o.classes(d))             # faults persist after the call


@pytest.mark.parametrize("net", CNV)
def test_tensor_core_and_popc_cores_agree(net):
    """Per-layer core selection (tcgen05 kind::i8 vs LOP3+POPC) must not change a single bit."""
    rt, o = runtime(net), Oracle(net)
    rec = synth_cifar(777, 41)
    want = o.cnv_scores(rec)
    for core in (2, 1, 0, 2):
        rt.set_core(core)
        cls, sc = rt.classify_cifar(rec, 10, want_scores=True)
        assert np.array_equal(sc, want), "core %d" % core


@pytest.mark.parametrize("net", LFC)
def test_lfc_cores_agree(net):
    rt, o = runtime(net), Oracle(net)
    px = synth_mnist(1000, 43)
    want = o.lfc_words(px)
    for core in (2, 1, 0):
        rt.set_core(core)
        _, words = rt.classify_mnist(px, 10, want_words=True)
        assert np.array_equal(words, want), "core %d" % core


def _n_gpus():
    import ctypes
    try:
        n = ctypes.c_int(0)
        rt = ctypes.CDLL("libcudart.so")
        return n.value if rt.cudaGetDeviceCount(ctypes.byref(n)) == 0 else 0
    except OSError:
        import torch
        return torch.cuda.device_count()


@pytest.mark.parametrize("net", ["cnvW1A1", "lfcW1A1"])
def test_two_gpu_shards_equal_one_gpu(net):
    """contiguous shards, replicated parameters, no collective; with faults every GPU applies the whole ordered list
    but runs only its own segments.  Needs two visible devices (skipped on a one-GPU box)."""
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    from bnn import native
    rt1 = native.Runtime(net, params_dir(net), devices=[0])
    rt2 = native.Runtime(net, params_dir(net), devices=[0, 1])
    n = 3001
    d = synth_cifar(n, 11) if net in CNV else synth_mnist(n, 11)
    assert np.array_equal(rt2.classify(d), rt1.classify(d))
    fl = rt1.schedule_faults(n, 30, word_size=1, target=-1, seed=5)
    assert np.array_equal(rt2.classify_with_fault_list(d, fl), rt1.classify_with_fault_list(d, fl))
    assert np.array_equal(rt2.classify(d), rt1.classify(d))         # the persistent faults are the same on every device
