"""The one-image call (bnn.py's inference(), `bnnb_classify_*` with n = 1) replays a captured CUDA graph from its second call on
(engine.cu run_single): the replay must pick up each call's own record, survive a reload of another parameter set into the same
engine, a change of number_class, batches in between, and agree with the ordinary path (BNN_B200_NO_GRAPH) and the oracle."""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle_py import NETS, Oracle, params_dir, synth_cifar, synth_mnist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))

pytestmark = pytest.mark.gpu


def one_by_one(rt, o, net, seed, count, ncls=10):
    cnv = net.startswith("cnv")
    imgs = synth_cifar(count, seed) if cnv else synth_mnist(count, seed)
    want_raw = o.cnv_scores(imgs) if cnv else o.lfc_words(imgs)
    for i in range(count):
        if cnv:
            cls, raw = rt.classify_cifar(imgs[i:i + 1], ncls, want_scores=True)
        else:
            cls, raw = rt.classify_mnist(imgs[i:i + 1], ncls, want_words=True)
        assert np.array_equal(raw, want_raw[i:i + 1]), (net, seed, i)
        batch = rt.classify_cifar(imgs, ncls) if cnv else rt.classify_mnist(imgs, ncls)  # the class rule, through the batch path
        assert int(cls[0]) == int(batch[i]), (net, seed, i)
        assert rt.usec_per_image > 0


@pytest.mark.parametrize("net", NETS)
def test_single_image_graph_replays_follow_their_inputs(net):
    from bnn import native
    rt, o = native.Runtime(net, params_dir(net), devices=[0]), Oracle(net)
    one_by_one(rt, o, net, 11, 6)
    one_by_one(rt, o, net, 12, 3, ncls=7)      # another number_class: another graph
    one_by_one(rt, o, net, 13, 3)
    rt.close()


def test_single_image_graph_after_reload_of_other_parameters():
    from bnn import native
    rt = native.Runtime("cnvW1A1", params_dir("cnvW1A1"), devices=[0])
    one_by_one(rt, Oracle("cnvW1A1"), "cnvW1A1", 21, 3)
    for ds in ("streetview", "road-signs"):
        rt.load_parameters(params_dir("cnvW1A1", ds))
        one_by_one(rt, Oracle("cnvW1A1", params_dir("cnvW1A1", ds)), "cnvW1A1", 22, 3)
    rt.close()


def test_single_image_graph_and_ordinary_path_agree():
    code = ("import sys, hashlib; sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2]);"
            "import numpy as np;"
            "from oracle_py import params_dir, synth_cifar; from bnn import native;"
            "rt = native.Runtime('cnvW1A2', params_dir('cnvW1A2'), devices=[0]); imgs = synth_cifar(8, 5); h = hashlib.sha256();\n"
            "for i in range(8):\n"
            "    cls, sc = rt.classify_cifar(imgs[i:i + 1], 10, want_scores=True); h.update(sc.tobytes() + cls.tobytes())\n"
            "print('hash', h.hexdigest())")
    got = {}
    for no_graph in (0, 1):
        env = dict(os.environ, BNN_B200_QUIET="1")
        if no_graph:
            env["BNN_B200_NO_GRAPH"] = "1"
        out = subprocess.run([sys.executable, "-c", code, os.path.join(ROOT, "bnn-pynq_b200"), os.path.join(ROOT, "tests")], env=env,
                             capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr[-2000:]
        got[no_graph] = out.stdout.split("hash")[1].split()[0]
    assert got[0] == got[1]


def test_single_image_graph_sees_injected_faults_and_memory_writes():
    """Faults and MemSet writes rewrite the device operands in place: a graph captured before them must compute with the new
    parameters (its nodes hold pointers, not values).  Single-image replays against the batch path on the same faulted engine,
    then against the oracle after the original parameters are loaded again."""
    from bnn import native
    net = "cnvW1A1"
    rt = native.Runtime(net, params_dir(net), devices=[0])
    imgs = synth_cifar(6, 31)
    for i in range(3):                                   # capture on clean parameters
        rt.classify_cifar(imgs[i:i + 1], 10)
    clean = rt.classify_cifar(imgs, 10, want_scores=True)[1]
    changed = False
    for layer, bit in ((8, 5), (7, 1234), (2, 77), (1, 3)):
        assert rt.weight_layer_bits(layer) > bit
        rt.inject_fault(0, layer, bit, 1)                # target 0: weights
        if rt.thresh_layer_bits(layer) > bit % 64:
            rt.inject_fault(1, layer, bit % 64, 1)       # target 1: thresholds
        batch = rt.classify_cifar(imgs, 10, want_scores=True)[1]
        changed = changed or not np.array_equal(batch, clean)
        for i in range(6):
            cls, sc = rt.classify_cifar(imgs[i:i + 1], 10, want_scores=True)
            assert np.array_equal(sc, batch[i:i + 1]), (layer, bit, i)
    assert changed
    rt.load_parameters(params_dir(net))
    o = Oracle(net)
    for i in range(6):
        cls, sc = rt.classify_cifar(imgs[i:i + 1], 10, want_scores=True)
        assert np.array_equal(sc, o.cnv_scores(imgs[i:i + 1]))
    rt.close()
