"""GPU parity at BASELINE.json's full sizes.  configs[1] and [2] (65 536 images): every image is distinct and the OpenMP oracle
classifies all of them.  configs[3] and [4] (32 768 ternary / 1 048 576 images with faults): the batch is drawn (with repetition,
in random order) from a few thousand distinct images the oracle HAS classified, so every result is still compared bit for bit;
the property behind that is the path's own: images are independent (DoCompute has no cross-image state, cnvW1A1/hw/top.cpp:180-236),
so a result may depend on the image and on the parameters in force when it is processed, never on its position, its neighbours or
the chunk / shard it lands in."""
import os
import sys

import numpy as np
import pytest

from oracle_py import Oracle, params_dir, synth_cifar, synth_mnist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
pytestmark = pytest.mark.gpu



def _runtime(net, all_devices=False):
    from bnn import native
    devs = [0]
    if all_devices:
        import torch
        devs = list(range(torch.cuda.device_count()))
    return native.Runtime(net, params_dir(net), devices=devs)


def _draw(unique, n, seed):
    idx = np.random.default_rng(seed).integers(0, unique.shape[0], n)
    idx[: unique.shape[0]] = np.arange(unique.shape[0])  # every unique image at least once
    return idx, np.ascontiguousarray(unique[idx])


@pytest.fixture(scope="module")
def cnv_distinct():
    """65 536 DISTINCT synthetic images (uniform bytes + the all-0 / all-255 images + 64 structured ones) and the oracle's scores and
    classes for every one of them: about two minutes of OpenMP oracle time on the GPU box's host cores, computed once per module."""
    rec = synth_cifar(65536, 101)
    img = rec[2:66, 1:].reshape(64, 3, 32, 32)
    for k in range(64):   # structured images: ramps of different slopes / constant blocks / one saturated channel
        if k % 4 == 0:
            img[k] = ((np.arange(32, dtype=np.int64)[None, None, :] * (k + 1)) % 256).astype(np.uint8)
        elif k % 4 == 1:
            img[k] = ((np.arange(32, dtype=np.int64)[None, :, None] * (k + 3)) % 256).astype(np.uint8)
        elif k % 4 == 2:
            img[k] = 0
            img[k, k % 3] = 255
        else:
            img[k] = 0
            img[k, :, : (k % 32) + 1, :] = 255
    o = Oracle("cnvW1A1")
    return rec, o.cnv_scores(rec), o.classes(rec)


def test_cnvW1A1_batch_65536_distinct_images_host_path(cnv_distinct):
    """configs[2]: CNV-W1A1, batch 65 536 on one GPU through the host-buffer C ABI: all 64 scores and the class of every one of
    65 536 distinct images against the oracle (no repetition draw)."""
    rec, want_sc, want_cls = cnv_distinct
    rt = _runtime("cnvW1A1")
    cls, sc = rt.classify_cifar(rec, 10, want_scores=True)
    assert np.array_equal(sc, want_sc)
    assert np.array_equal(cls, want_cls)
    # idempotence: the same images in another order
    cls2 = rt.classify_cifar(rec[::-1], 10)
    assert np.array_equal(cls2[::-1], cls)


def test_cnvW1A1_batch_65536_distinct_images_device_resident_path(cnv_distinct):
    """The path bench.py's `value` times - one 65 536-image device-resident chunk (bnnb_prep_device + bnnb_infer_device) - checked
    directly against the oracle, scores and classes of every image."""
    import torch
    rec, want_sc, want_cls = cnv_distinct
    rt = _runtime("cnvW1A1")
    cls, sc = rt.classify_tensor(torch.from_numpy(rec).cuda(), 10, want_scores=True)
    torch.cuda.synchronize()
    rt.sync()   # also reports a kernel that gave up on an mbarrier
    assert np.array_equal(sc.cpu().numpy(), want_sc)
    assert np.array_equal(cls.cpu().numpy(), want_cls)


@pytest.mark.parametrize("net", ["cnvW1A2", "cnvW2A2"])
def test_ternary_cnv_batch_32768_scores(net):
    """configs[3]: the 2-bit networks at a full device chunk (32 768 images): 4 096 distinct images classified by the oracle, the rest
    of the chunk drawn from them with repetition (images are independent: a result depends on the image and the parameters only)."""
    rt, o = _runtime(net), Oracle(net)
    uniq = synth_cifar(4096, 103)
    want = o.cnv_scores(uniq)
    idx, batch = _draw(uniq, 32768, 3)
    _, sc = rt.classify_cifar(batch, 10, want_scores=True)
    assert np.array_equal(sc, want[idx])


def test_lfcW1A2_batch_65536_distinct_words():
    """configs[1]: LFC-W1A2, batch 65 536 on one GPU: the 64-bit output word and class of every one of 65 536 distinct images, through
    the host-buffer path and through the device-resident one."""
    import torch
    rt, o = _runtime("lfcW1A2"), Oracle("lfcW1A2")
    px = synth_mnist(65536, 102)
    want, want_cls = o.lfc_words(px), o.classes(px)
    cls, words = rt.classify_mnist(px, 10, want_words=True)
    assert np.array_equal(words, want)
    assert np.array_equal(cls, want_cls)
    dcls, dwords = rt.classify_tensor(torch.from_numpy(px).cuda(), 10, want_scores=True)
    torch.cuda.synchronize()
    rt.sync()
    assert np.array_equal(dwords.cpu().numpy().view(np.uint64), want)
    assert np.array_equal(dcls.cpu().numpy(), want_cls)


def test_cnvW1A1_one_million_images_all_devices_equal_one_device_with_and_without_faults():
    """configs[4]: CNV-W1A1, 1 048 576 images sharded over every visible GPU by the engine itself (one process, one feeder thread per
    device), with and without a 1 024-fault schedule: identical to the one-device run, and the unfaulted classes equal the oracle's
    (the batch is drawn from 2 048 distinct images the oracle has classified; the faulted run is pinned to the oracle at 262 144
    images by the next test)."""
    n = 1 << 20
    o = Oracle("cnvW1A1")
    uniq = synth_cifar(2048, 105)
    idx, batch = _draw(uniq, n, 5)
    rt1, rtg = _runtime("cnvW1A1"), _runtime("cnvW1A1", all_devices=True)
    want = o.classes(uniq)[idx]
    got1 = rt1.classify_cifar(batch, 10)
    assert np.array_equal(got1, want)
    assert np.array_equal(rtg.classify_cifar(batch, 10), want)
    fl = rt1.schedule_faults(n, 1024, word_size=1, target=-1, seed=11)
    assert fl.shape[0] > 1000
    f1 = rt1.classify_with_fault_list(batch, fl)
    fg = rtg.classify_with_fault_list(batch, fl)
    assert np.array_equal(f1, fg)
    assert not np.array_equal(f1, want)                        # 1 024 upsets do change results
    first = int(fl[:, 0].min())
    assert np.array_equal(f1[:first], want[:first])            # nothing changes before the first fault time
    assert np.array_equal(rtg.classify_cifar(batch[:4096], 10), rt1.classify_cifar(batch[:4096], 10))  # the same faults persist on every device


def test_cnvW1A1_large_batch_with_scheduled_faults_all_devices():
    """configs[4] (256 Ki images instead of 1 Mi to bound host memory): a fault list spread over the batch, the batch sharded over
    every visible GPU.  Image i must be classified under exactly the faults with time <= i (faults.h:138-146), whatever chunk or
    shard it falls into: the oracle classifies the unique images once per cumulative fault state."""
    n = 262144
    rt, o = _runtime("cnvW1A1", all_devices=True), Oracle("cnvW1A1")
    uniq = synth_cifar(256, 104)
    idx, batch = _draw(uniq, n, 4)
    fl = rt.schedule_faults(n, 12, word_size=1, target=-1, seed=9)
    times = np.array([0, 1, 8191, 8192, 65535, 65536, 100000, 131071, 131072, 200001, 200001, n - 1], np.uint32)[: fl.shape[0]]
    assert times.size >= 8
    fl[:, 0] = times
    got = rt.classify_with_fault_list(batch, fl)
    want = np.empty(n, np.int32)
    bounds = sorted(set(int(t) for t in times)) + [n]
    state_cls = o.classes(uniq)
    if bounds[0] > 0:
        want[: bounds[0]] = state_cls[idx[: bounds[0]]]
    for a, b in zip(bounds[:-1], bounds[1:]):
        now = fl[fl[:, 0] == a].copy()
        now[:, 0] = 0
        state_cls = o.classify_with_faults(uniq, now)  # cumulative: the oracle keeps its faults, like the reference
        want[a:b] = state_cls[idx[a:b]]
    assert np.array_equal(got, want)
