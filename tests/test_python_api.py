"""GPU tests of the bnn.py mirror: the six cffi symbols of lib/python_cuda-<net>.so driven exactly as bnn/bnn.py drives the
reference libraries (known answers of tests/test_qnn.py:48-133 and the golden vectors of the reference runtime)."""
import os
import sys
import tempfile

import numpy as np
import pytest

from oracle_py import GOLDEN, Oracle, mnist_idx_bytes, synth_cifar, synth_mnist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
pytestmark = pytest.mark.gpu
os.environ.setdefault("BNN_B200_QUIET", "1")


def _bnn():
    import bnn
    bnn.bnn.set_param_dir(os.path.join(GOLDEN, "params"))
    return bnn


def _tmp(data):
    f = tempfile.NamedTemporaryFile(delete=False)
    f.write(data)
    f.close()
    return f.name


@pytest.mark.parametrize("net", ["cnvW1A1", "cnvW1A2", "cnvW2A2"])
def test_cnv_classifier(net):
    bnn = _bnn()
    g = np.load(os.path.join(GOLDEN, "ref_%s.npz" % net))
    c = bnn.CnvClassifier(net, "cifar10", bnn.RUNTIME_CUDA)
    assert len(c.classes) == 10 and c.class_name(4) == "Deer"
    deer = os.path.join(GOLDEN, "images", "deer.bin")
    assert c.classify_cifar(deer) == 4 == g["deer_class_single"][0]
    assert np.array_equal(c.classify_cifar_details(deer), g["deer_single_scores"])
    assert c.usecPerImage > 0
    path = _tmp(synth_cifar(32, 0).tobytes())
    assert np.array_equal(c.classify_cifars(path), g["synth_classes"])
    assert np.array_equal(c.classify_cifars_details(path).reshape(32, 10), g["synth_scores"])
    clean = c.classify_cifars(path)
    assert np.array_equal(c.classify_cifars_with_faults(path, 0, 1, -1), clean)            # no faults: same answer
    faulty = c.classify_cifars_with_faults(path, 3000, 1, 1, [1, 2, 3])                    # threshold faults in layers 1-3
    assert faulty.shape == clean.shape and (faulty != clean).any()
    c2 = bnn.CnvClassifier(net, "cifar10", bnn.RUNTIME_CUDA)                               # a new classifier reloads parameters
    assert np.array_equal(c2.classify_cifars(path), g["synth_classes"])
    os.unlink(path)


def test_cnv_classify_image_roundtrip():
    pytest.importorskip("PIL")
    from PIL import Image
    bnn = _bnn()
    c = bnn.CnvClassifier("cnvW1A1", "cifar10", bnn.RUNTIME_CUDA)
    rec = np.frombuffer(open(os.path.join(GOLDEN, "images", "deer.bin"), "rb").read(), np.uint8)
    img = Image.fromarray(rec[1:].reshape(3, 32, 32).transpose(1, 2, 0).copy(), "RGB")     # already 32x32: image_to_cifar is exact
    assert c.classify_image(img) == 4
    assert list(c.classify_images([img, img])) == [4, 4]


@pytest.mark.parametrize("net", ["lfcW1A1", "lfcW1A2"])
def test_lfc_classifier(net):
    bnn = _bnn()
    g = np.load(os.path.join(GOLDEN, "ref_%s.npz" % net))
    c = bnn.LfcClassifier(net, "mnist", bnn.RUNTIME_CUDA)
    assert c.classify_mnist(os.path.join(GOLDEN, "images", "3.image-idx3-ubyte")) == 3
    path = _tmp(mnist_idx_bytes(synth_mnist(256, 0)))
    assert np.array_equal(c.classify_mnists(path), g["synth_classes"])
    assert np.array_equal(c.classify_mnists_with_faults(path, 0, 1, -1), g["synth_classes"])
    os.unlink(path)


def test_other_parameter_sets_short_files():
    """streetview / road-signs (cnvW1A1) and chars_merged (lfcW1A1, 47 classes, short layer-3 files): loader zero-fill."""
    bnn = _bnn()
    from oracle_py import params_dir
    rec = synth_cifar(64, 8)
    path = _tmp(rec.tobytes())
    for ds in ("streetview", "road-signs"):
        c = bnn.CnvClassifier("cnvW1A1", ds, bnn.RUNTIME_CUDA)
        o = Oracle("cnvW1A1", params_dir("cnvW1A1", ds))
        assert np.array_equal(c.classify_cifars(path), o.classes(rec, len(c.classes)))
    os.unlink(path)
    px = synth_mnist(300, 9)
    path = _tmp(mnist_idx_bytes(px))
    c = bnn.LfcClassifier("lfcW1A1", "chars_merged", bnn.RUNTIME_CUDA)
    o = Oracle("lfcW1A1", params_dir("lfcW1A1", "chars_merged"))
    assert len(c.classes) == 47
    assert np.array_equal(c.classify_mnists(path), o.classes(px, 47))
    os.unlink(path)
