"""CPU tests: the C-ABI libraries load and export every symbol include/bnn_b200.h declares; the host-side parameter mirror
(loader, MemRead/MemSet, fault map) agrees with the oracle and with the reference read-backs; no compute without a GPU."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from oracle_py import GOLDEN, NETS, Oracle, params_dir

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "bnn-pynq_b200")
sys.path.insert(0, PKG)
SIX = ["load_parameters", "inference", "inference_multiple", "inference_multiple_with_faults", "free_results", "deinit"]
BUFFER_TWINS = ["inference_buffer", "inference_buffer_with_faults"]   # in-memory extension (include/bnn_b200.h section 1)


@pytest.fixture(scope="session", autouse=True)
def built():
    if not os.path.exists(os.path.join(PKG, "lib", "libbnn_b200.so")):
        subprocess.check_call(["make", "-C", PKG, "-j8"])


def declared():
    src = open(os.path.join(ROOT, "include", "bnn_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bnnb_\w+)\s*\(", src)))


def test_core_library_exports_every_declared_symbol():
    lib = C.CDLL(os.path.join(PKG, "lib", "libbnn_b200.so"))
    names = declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), n


@pytest.mark.parametrize("net", NETS)
def test_per_network_library_exports_the_six_bnn_py_symbols(net):
    lib = C.CDLL(os.path.join(PKG, "lib", "python_cuda-%s.so" % net))
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "bnn_b200.h")).read(), flags=re.S)
    section1 = src[:src.index("typedef struct bnnb_ctx")]
    names = sorted(set(re.findall(r"\b(\w+)\s*\(", section1)))
    assert names == sorted(SIX + BUFFER_TWINS)                          # the header declares exactly these for the per-network objects
    for n in names:
        assert hasattr(lib, n), n


def test_python_mirror_surface():
    import bnn
    assert bnn.RUNTIME_CUDA == "python_cuda"
    for name in ("classify_image", "classify_images", "classify_cifar", "classify_cifars", "classify_cifar_details", "classify_cifars_details",
                 "classify_path", "classify_paths", "classify_images_with_faults", "classify_cifars_with_faults", "class_name"):
        assert hasattr(bnn.CnvClassifier, name)
    for name in ("classify_mnist", "classify_mnists", "classify_mnists_with_faults", "class_name"):
        assert hasattr(bnn.LfcClassifier, name)
    bnn.bnn.set_param_dir(os.path.join(GOLDEN, "params"))
    assert "cifar10" in bnn.available_params("cnvW1A1") and "mnist" in bnn.available_params("lfcW1A2")


@pytest.mark.parametrize("net", NETS)
def test_host_mirror_matches_oracle_and_reference_readbacks(net):
    from bnn import native
    import torch
    rt = native.Runtime(net)
    if torch.cuda.is_available():
        rt.load_parameters(params_dir(net))
    else:
        with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):
            rt.load_parameters(params_dir(net))  # the mirror is filled, the device upload fails loudly
    o = Oracle(net)
    g = np.load(os.path.join(GOLDEN, "ref_%s.npz" % net))
    rng = np.random.default_rng(5)
    for l in range(o.nlayers):
        li = o.layer_info(l)
        assert rt.weight_layer_bits(l) == o.weight_layer_bits(l) and rt.thresh_layer_bits(l) == o.thresh_layer_bits(l)
        for _ in range(50):
            pe, ind = int(rng.integers(li["pe"])), int(rng.integers(li["wmem"]))
            assert rt.mem_read(2 * l, pe, ind, 0) == o.mem_read(2 * l, pe, ind, 0)
            ti, th = int(rng.integers(li["tmem"])), int(rng.integers(max(li["nthr"], 1)))
            assert rt.mem_read(2 * l + 1, pe, ti, th) == o.mem_read(2 * l + 1, pe, ti, th)
    for t, l, b in g["fault_list"]:
        try:
            rt.inject_fault(int(t), int(l), int(b), 1)
        except RuntimeError:
            assert not torch.cuda.is_available()
    reads = []
    for l in range(o.nlayers):
        li = o.layer_info(l)
        reads += [rt.mem_read(2 * l, 0, 0, 0), rt.mem_read(2 * l + 1, li["pe"] - 1, li["tmem"] - 1, 0)]
    assert np.array_equal(np.array(reads, np.uint64), g["fault_reads"])


def test_random_fault_schedule_follows_the_reference_distributions():
    from bnn import native
    rt = native.Runtime("cnvW1A1")
    try:
        rt.load_parameters(params_dir("cnvW1A1"))
    except RuntimeError:
        pass
    n = 1000
    fl = rt.schedule_faults(n, 4000, word_size=8, target=-1, seed=3)
    assert fl.shape == (4000, 5) and (np.diff(fl[:, 0].astype(np.int64)) >= 0).all() and fl[:, 0].max() < n
    assert set(fl[:, 4]) == {1}                                   # the bool collapse of word_size (faults.h:42 vs :55)
    wb = sum(rt.weight_layer_bits(l) for l in range(9))
    tb = sum(rt.thresh_layer_bits(l) for l in range(9))
    frac_w = (fl[:, 1] == 1).mean()
    assert abs(frac_w - wb / (wb + tb)) < 0.02                    # weights vs thresholds in proportion to their bit counts
    for row in fl:
        bits = rt.weight_layer_bits(int(row[2])) if row[1] == 1 else rt.thresh_layer_bits(int(row[2]))
        assert row[3] < bits
    assert not ((fl[:, 1] == 2) & (fl[:, 2] == 8)).any()          # layer 8 has no threshold bits
    only = rt.schedule_faults(n, 300, 1, 0, [1, 3], seed=4)
    assert set(only[:, 1]) == {1} and set(only[:, 2]) <= {1, 3}
    assert np.array_equal(rt.schedule_faults(n, 50, 1, -1, seed=9), rt.schedule_faults(n, 50, 1, -1, seed=9))
    assert rt.schedule_faults(n, 0).shape[0] == 0


def test_compute_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from bnn import native
    rt = native.Runtime("lfcW1A1")
    with pytest.raises(RuntimeError):
        rt.classify_mnist(np.zeros((1, 784), np.uint8))


def test_bind_host_to_device_never_raises():
    """native.bind_host_to_device pins a rank to its GPU's CPUs when NVML knows them and is a no-op (None) otherwise; run in a
    child process so that this process keeps its affinity."""
    import subprocess
    code = ("import sys, os; sys.path.insert(0, %r); from bnn import native; before = os.sched_getaffinity(0); "
            "r = native.bind_host_to_device(0); after = os.sched_getaffinity(0); "
            "assert r is None or (len(r) > 0 and set(r) == set(after) and set(after) <= set(before)); "
            "assert r is not None or after == before; print('ok')") % os.path.join(ROOT, "bnn-pynq_b200")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr


def test_hardened_variants_are_named_and_refused_explicitly():
    """The fork's -TMR / -interleaved networks (bnn/bnn.py:41-53) are exported names but an explicit NotImplementedError on the CUDA
    runtime - not a misleading 'library not built' message."""
    sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
    import bnn
    assert bnn.NETWORK_CNVW1A1_TMR == "cnvW1A1-TMR" and bnn.NETWORK_LFCW1A2_INTERLEAVED == "lfcW1A2-interleaved"
    with pytest.raises(NotImplementedError, match="hardened variant"):
        bnn.PynqBNN(bnn.RUNTIME_CUDA, bnn.NETWORK_CNVW1A1_TMR)
