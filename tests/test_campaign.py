"""Fault-campaign drivers (bnn.util / bnn.faults, the counterparts of bnn/util/util.py and bnn/faults/faults.py of BNN-PYNQ).
CPU part: the helpers on synthetic files.  GPU part: a small campaign through the CUDA runtime — zero flips reproduces the control
run, faults are discarded between runs, the result files carry the reference's keys."""
import json
import os
import sys

import numpy as np
import pytest

from oracle_py import GOLDEN, mnist_idx_bytes, synth_cifar, synth_mnist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
os.environ.setdefault("BNN_B200_QUIET", "1")


def _util():
    from bnn import util
    return util


def test_calculate_accuracy():
    u = _util()
    assert u.calculate_accuracy([1, 2, 3, 4], [1, 2, 0, 4]) == 75.0
    assert u.calculate_accuracy(np.array([1, 2, 3, 4, 9, 9]), [1, 2, 3, 4]) == 100.0        # results may be longer than the labels
    assert u.calculate_accuracy([], []) == 0.0


def test_dict_of_dicts_merge():
    u = _util()
    a = {"network": "cnvW1A1", "results": {"any bit": [1.0, 2.0]}, "control": 80.0}
    b = {"network": "other", "results": {"weight bit": [3.0]}, "layers": [1]}
    m = u.dict_of_dicts_merge(a, b)
    assert m == {"network": "cnvW1A1", "control": 80.0, "layers": [1], "results": {"any bit": [1.0, 2.0], "weight bit": [3.0]}}
    m["results"]["any bit"].append(5.0)
    assert a["results"]["any bit"] == [1.0, 2.0]                                            # deep copies, inputs untouched


def test_write_dict_to_file_roundtrip(tmp_path):
    u = _util()
    d = {"a": {"b": [1, 2.5, "x"]}, "c": None}
    p = tmp_path / "x" / "y" / "out.json"
    u.write_dict_to_file(str(p), d)
    assert json.loads(p.read_text()) == d
    assert json.loads(u.dict_to_str(d)) == d


def test_load_cifar10_testset(tmp_path):
    u = _util()
    rec = synth_cifar(40, 3)
    # the loader caps at 10 000 and reads test_batch.bin; asking for fewer writes test_batch_<n>.bin once
    (tmp_path / "test_batch.bin").write_bytes(rec.tobytes())
    path, labels = u.load_cifar10_testset(str(tmp_path), 16)
    assert path.endswith("test_batch_16.bin") and os.path.getsize(path) == 16 * 3073
    assert labels == [int(x) for x in rec[:16, 0]]
    assert np.array_equal(np.fromfile(path, np.uint8), rec[:16].reshape(-1))
    path2, labels2 = u.load_cifar10_testset(str(tmp_path), 16)                              # second call reuses the file
    assert path2 == path and labels2 == labels


def test_load_mnist_testset(tmp_path):
    u = _util()
    lbl = np.arange(50, dtype=np.uint8) % 10
    (tmp_path / "t10k-labels-idx1-ubyte").write_bytes(bytes([0, 0, 8, 1, 0, 0, 0, 50]) + lbl.tobytes())
    (tmp_path / "t10k-images-idx3-ubyte").write_bytes(mnist_idx_bytes(synth_mnist(50, 1)))
    path, labels = u.load_mnist_testset(str(tmp_path), 20)
    assert path.endswith("t10k-images-idx3-ubyte") and labels == [int(x) for x in lbl[:20]]


def test_load_gtsrb_testset(tmp_path):
    pytest.importorskip("PIL")
    from PIL import Image
    u = _util()
    rows = ["Filename;Width;Height;Roi.X1;Roi.Y1;Roi.X2;Roi.Y2;ClassId"]
    for i in range(5):
        Image.fromarray(np.full((8, 8, 3), i * 10, np.uint8), "RGB").save(str(tmp_path / ("%05d.ppm" % i)))
        rows.append("%05d.ppm;8;8;0;0;7;7;%d" % (i, 40 - i))
    (tmp_path / "gt.csv").write_text("\n".join(rows) + "\n")
    images, labels = u.load_gtsrb_testset(str(tmp_path / "gt.csv"), str(tmp_path), 3)
    assert labels == [40, 39, 38] and len(images) == 3 and images[2].getpixel((0, 0)) == (20, 20, 20)


def test_test_type_names_and_stats():
    from bnn import faults
    T = faults.NetworkTest.TestType
    assert [t.name for t in (T.any_bit(), T.any_word(), T.weight_bit(), T.weight_word(), T.threshold_bit(), T.threshold_word())] == \
        ["any bit", "any word", "weight bit", "weight word", "threshold bit", "threshold word"]
    assert T.any_word().word_size == 8 and T.weight_word(4).word_size == 4 and T.threshold_bit().target_type == 1
    nt = faults.NetworkTest(fault_test=None)
    s = nt._stats({"control": 80.0, "results": {"any bit": [80.0, 70.0, 60.0], "weight bit": [80.0, 80.0]}})
    e = s["results"]["any bit"]
    assert e["runs"] == {"all": [80.0, 70.0, 60.0], "effective": [70.0, 60.0]} and e["effective count"] == 2
    assert e["min accuracy"] == 60.0 and e["max accuracy"] == 80.0 and e["avg accuracy"] == 70.0
    assert e["avg effective accuracy"] == 65.0 and e["accuracy delta"] == 10.0 and e["effective accuracy delta"] == 15.0
    w = s["results"]["weight bit"]
    assert w["effective count"] == 0 and w["avg accuracy"] == 80.0 and "accuracy delta" not in w


# ---------------------------------------------------------------- GPU: campaigns through the CUDA runtime
def _bnn():
    import bnn
    bnn.bnn.set_param_dir(os.path.join(GOLDEN, "params"))
    return bnn


@pytest.mark.gpu
def test_cnv_campaign(tmp_path):
    bnn = _bnn()
    from bnn import faults, util
    rec = synth_cifar(96, 21)
    (tmp_path / "test_batch.bin").write_bytes(rec.tobytes())
    path, _ = util.load_cifar10_testset(str(tmp_path), 64)
    clean = [int(x) for x in bnn.CnvClassifier("cnvW1A1", "cifar10", bnn.RUNTIME_CUDA).classify_cifars(path)]
    labels = list(clean)
    labels[0] = (labels[0] + 1) % 10                                                        # control accuracy 63/64
    ft = faults.CNVFaultTest.CIFARTest("cnvW1A1", path, labels, quiet=True)
    results, times, acc = ft.run_test(num_runs=2, num_flips=0)
    assert results == [clean, clean] and acc == [63 * 100 / 64] * 2 and all(t > 0 for t in times)
    # heavy threshold upsets in the first layers change answers; the next zero-flip run is clean again (faults discarded)
    r_faulty, _, acc_faulty = ft.run_test(2, 4000, 8, faults.FaultTest.TargetType.thresholds(), [0, 1, 2])
    assert len(r_faulty) == 2 and all(len(r) == 64 for r in r_faulty) and min(acc_faulty) < acc[0]
    assert r_faulty[0] != r_faulty[1]                                                       # runs draw different faults
    assert ft.run_test(1, 0)[0][0] == clean
    nt = faults.NetworkTest(ft)
    T = faults.NetworkTest.TestType
    written = nt.test_network(str(tmp_path / "out"), 2, [1, 2000], [T.any_bit(), T.threshold_word()], [1, 2])
    assert nt.control == 63 * 100 / 64 and len(written) == 2
    for w, flips in zip(written, (1, 2000)):
        assert w.endswith(os.path.join("cnvW1A1", "cifar10", "%dflips" % flips, "cnvW1A1_cifar10_stats_layer[1, 2].json"))
        s = json.load(open(w))
        assert s["network"] == "cnvW1A1" and s["dataset"] == "cifar10" and s["run count"] == 2 and s["flips"] == flips
        assert s["layers"] == [1, 2] and s["control"] == nt.control and set(s["results"]) == {"any bit", "threshold word"}
        for e in s["results"].values():
            assert len(e["runs"]["all"]) == 2 and e["effective count"] == len(e["runs"]["effective"])
        raw = json.load(open(os.path.join(os.path.dirname(w), "temp", "cnvW1A1_results_threshold-word.json")))
        assert raw["results"]["threshold word"] == s["results"]["threshold word"]["runs"]["all"]
    assert json.load(open(written[1]))["results"]["threshold word"]["effective count"] == 2


@pytest.mark.gpu
def test_lfc_campaign_comprehensive(tmp_path):
    bnn = _bnn()
    from bnn import faults
    px = synth_mnist(128, 5)
    p = tmp_path / "t10k-images-idx3-ubyte"
    p.write_bytes(mnist_idx_bytes(px))
    clean = [int(x) for x in bnn.LfcClassifier("lfcW1A1", "mnist", bnn.RUNTIME_CUDA).classify_mnists(str(p))]
    ft = faults.LFCFaultTest.MNISTTest("lfcW1A1", str(p), clean[:100], quiet=True)           # labels shorter than the file
    nt = faults.NetworkTest(ft)
    written = nt.comprehensive_test(str(tmp_path / "out"), 2, [500])
    assert nt.control == 100.0 and len(written) == 1 and written[0].endswith("lfcW1A1_mnist_stats.json")
    s = json.load(open(written[0]))
    assert list(s["results"]) == ["any bit", "any word", "weight bit", "weight word", "threshold bit", "threshold word"]
    assert s["layers"] == [] and all(len(e["runs"]["all"]) == 2 for e in s["results"].values())


@pytest.mark.gpu
def test_image_list_campaign():
    """Datasets other than cifar10 / mnist go through classify_images_with_faults (PIL images), as the reference's SVHN / GTSRB
    campaigns do (bnn/faults/faults.py:66-68)."""
    pytest.importorskip("PIL")
    from PIL import Image
    bnn = _bnn()
    from bnn import faults
    rec = synth_cifar(24, 9)
    imgs = [Image.fromarray(r[1:].reshape(3, 32, 32).transpose(1, 2, 0).copy(), "RGB") for r in rec]
    clean = [int(x) for x in bnn.CnvClassifier("cnvW1A1", "streetview", bnn.RUNTIME_CUDA).classify_images(imgs)]
    ft = faults.CNVFaultTest.SVHNTest("cnvW1A1", imgs, clean, quiet=True)
    results, _, acc = ft.run_test(1, 0)
    assert results == [clean] and acc == [100.0]
