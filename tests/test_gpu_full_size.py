"""GPU parity at BASELINE.json's full sizes (configs[1], [2] and - scaled to 256 Ki images - [4]), where the CPU oracle cannot be run
on every image: the batch is drawn (with repetition, in random order) from a small set of unique images the oracle HAS classified,
so every one of the 65 536+ results is still compared bit for bit.  The property behind it is the path's own: images are
independent (DoCompute has no cross-image state, cnvW1A1/hw/top.cpp:180-236), so a result may depend on the image and on the
parameters in force when it is processed, never on its position, its neighbours or the chunk / shard it lands in."""
import os
import sys

import numpy as np
import pytest

from oracle_py import Oracle, params_dir, synth_cifar, synth_mnist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
pytestmark = pytest.mark.gpu

UNIQUE = 768


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


def test_cnvW1A1_batch_65536_scores():
    """configs[2]: CNV-W1A1, batch 65 536 on one GPU: all 64 scores and the class of every image."""
    rt, o = _runtime("cnvW1A1"), Oracle("cnvW1A1")
    uniq = synth_cifar(UNIQUE, 101)
    uniq[0, 1:] = 0      # all-0 and all-255 images (the arg-max floor and the saturating input quantiser)
    uniq[1, 1:] = 255
    want_sc, want_cls = o.cnv_scores(uniq), o.classes(uniq)
    idx, batch = _draw(uniq, 65536, 1)
    cls, sc = rt.classify_cifar(batch, 10, want_scores=True)
    assert np.array_equal(sc, want_sc[idx])
    assert np.array_equal(cls, want_cls[idx])
    # idempotence: the same call again, and the same images in another order
    cls2 = rt.classify_cifar(batch[::-1], 10)
    assert np.array_equal(cls2[::-1], cls)


@pytest.mark.parametrize("net", ["cnvW1A2", "cnvW2A2"])
def test_ternary_cnv_batch_32768_scores(net):
    """configs[3]: the 2-bit networks at a full device chunk (32 768 images)."""
    rt, o = _runtime(net), Oracle(net)
    uniq = synth_cifar(256, 103)
    want = o.cnv_scores(uniq)
    idx, batch = _draw(uniq, 32768, 3)
    _, sc = rt.classify_cifar(batch, 10, want_scores=True)
    assert np.array_equal(sc, want[idx])


def test_lfcW1A2_batch_65536_words():
    """configs[1]: LFC-W1A2, batch 65 536 on one GPU: the 64-bit output word of every image."""
    rt, o = _runtime("lfcW1A2"), Oracle("lfcW1A2")
    uniq = synth_mnist(2048, 102)
    want = o.lfc_words(uniq)
    idx, batch = _draw(uniq, 65536, 2)
    cls, words = rt.classify_mnist(batch, 10, want_words=True)
    assert np.array_equal(words, want[idx])
    assert np.array_equal(cls, o.classes(uniq)[idx])


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
