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


# ---------------------------------------------------------------- in-memory ingestion
@pytest.mark.parametrize("net", ["cnvW1A1", "cnvW2A2"])
def test_cnv_in_memory_matches_path_based(net):
    bnn = _bnn()
    g = np.load(os.path.join(GOLDEN, "ref_%s.npz" % net))
    c = bnn.CnvClassifier(net, "cifar10", bnn.RUNTIME_CUDA)
    rec = synth_cifar(32, 0)
    assert np.array_equal(c.classify_records(rec), g["synth_classes"])
    assert np.array_equal(c.classify_records(rec, details=True).reshape(32, 10), g["synth_scores"])
    assert np.array_equal(c.classify_records(rec[:, 1:]), g["synth_classes"])              # 3072-byte rows: label byte optional
    assert np.array_equal(c.classify_records(rec.reshape(-1)), g["synth_classes"])         # flat record stream, as read from a file
    assert c.usecPerImage > 0
    assert c.classify_records(rec[:0]).shape == (0,)                                       # empty batch
    assert np.array_equal(c.classify_records_with_faults(rec, 0, 1, -1), g["synth_classes"])
    faulty = c.classify_records_with_faults(rec, 3000, 1, 1, [1, 2, 3])
    assert faulty.shape == (32,) and (faulty != g["synth_classes"]).any()
    with pytest.raises(ValueError):
        c.classify_records(np.zeros((2, 100), np.uint8))


def test_cnv_classify_images_in_memory_equals_temp_file_route():
    pytest.importorskip("PIL")
    from PIL import Image
    bnn = _bnn()
    c = bnn.CnvClassifier("cnvW1A1", "cifar10", bnn.RUNTIME_CUDA)
    rng = np.random.default_rng(4)
    imgs = [Image.fromarray(rng.integers(0, 256, (h, w, 3), dtype=np.uint8), "RGB") for h, w in ((32, 32), (48, 64), (20, 31), (100, 40))]
    rec = c.images_to_records([im.copy() for im in imgs])
    assert rec.shape == (4, 3073) and (rec[:, 0] == 1).all()                               # image_to_cifar writes label byte 1
    via_file = c._with_tmp([im.copy() for im in imgs], c.bnn.inference_multiple)           # the reference's route (bnn/bnn.py:280-287)
    assert np.array_equal(c.classify_images([im.copy() for im in imgs]), via_file)
    det_file = c._with_tmp([im.copy() for im in imgs], c.bnn.inference_multiple_detail)
    assert np.array_equal(c.classify_images_details([im.copy() for im in imgs]), det_file)


def test_lfc_in_memory_matches_path_based():
    bnn = _bnn()
    g = np.load(os.path.join(GOLDEN, "ref_lfcW1A1.npz"))
    c = bnn.LfcClassifier("lfcW1A1", "mnist", bnn.RUNTIME_CUDA)
    px = synth_mnist(256, 0)
    assert np.array_equal(c.classify_pixels(px), g["synth_classes"])
    assert np.array_equal(c.classify_pixels(px.reshape(256, 28, 28)), g["synth_classes"])
    assert np.array_equal(c.classify_pixels_with_faults(px, 0, 1, -1), g["synth_classes"])
    assert c.classify_pixels(px[:1]).shape == (1,)


@pytest.mark.parametrize("net", ["cnvW1A1", "cnvW1A2", "lfcW1A1", "lfcW1A2"])
def test_device_tensor_entry_point(net):
    """CUDA tensors / DLPack in, CUDA tensors out: same classes, scores and words as the host-buffer call."""
    import torch
    from bnn import native
    from oracle_py import params_dir
    rt = native.Runtime(net, params_dir(net), devices=[0])
    if net.startswith("cnv"):
        data = synth_cifar(300, 12)
        cls, sc = rt.classify_cifar(data, 10, want_scores=True)
    else:
        data = synth_mnist(300, 12)
        cls, sc = rt.classify_mnist(data, 10, want_words=True)
    t = torch.from_numpy(data).cuda()
    d_cls, d_sc = rt.classify_tensor(t, 10, want_scores=True)
    assert d_cls.is_cuda and d_cls.dtype == torch.int32
    assert np.array_equal(d_cls.cpu().numpy(), cls)
    assert np.array_equal(d_sc.cpu().numpy().view(sc.dtype).reshape(sc.shape), sc)

    class Capsule:                                                                          # a foreign DLPack exporter
        def __init__(self, t):
            self.t = t

        def __dlpack__(self, **kw):
            return self.t.__dlpack__(**kw)

        def __dlpack_device__(self):
            return self.t.__dlpack_device__()
    assert np.array_equal(rt.classify_tensor(Capsule(t)).cpu().numpy(), cls)
    assert rt.classify_tensor(t[:0]).shape == (0,)
    if not net.startswith("cnv"):                                                           # all 64 mask widths of the LFC decode
        for nc in (1, 7, 47, 64):
            assert np.array_equal(rt.classify_tensor(t, nc).cpu().numpy(), rt.classify_mnist(data, nc))
    with pytest.raises(ValueError):
        rt.classify_tensor(torch.from_numpy(data))                                          # host tensor: refused, no silent copy
