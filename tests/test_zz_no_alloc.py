"""bnnb_infer_device_static - the entry point that allocates nothing (static 32 MiB device scratch; include/bnn_b200.h) - against
bnnb_infer_device through the C ABI, bit for bit (integer path: tolerance 0, as in every parity test here), on two calls in a row
with different inputs, so that the second call runs on the first one's leftovers in the shared scratch.  Collected last on purpose;
it passes alone in a fresh process and after the other tests."""
import os
import sys

import numpy as np
import pytest

from oracle_py import params_dir, synth_cifar, synth_mnist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
pytestmark = pytest.mark.gpu

# (network, images of call 1, images of call 2): within the static limit, ragged against the 128-image tiles
CASES = [("lfcW1A1", 3000, 32768), ("lfcW1A2", 1025, 257), ("cnvW1A1", 1792, 77), ("cnvW1A2", 200, 1152), ("cnvW2A2", 131, 64)]
LIMIT = {"lfcW1A1": 32768, "lfcW1A2": 32768, "cnvW1A1": 1792, "cnvW1A2": 1792, "cnvW2A2": 1792}


@pytest.mark.parametrize("net,n1,n2", CASES)
def test_static_scratch_entry_point_equals_infer_device(net, n1, n2):
    import torch
    from bnn import native
    rt = native.Runtime(net, params_dir(net), devices=[0])
    cnv = net.startswith("cnv")
    dev = torch.device("cuda", 0)

    def run(n, seed, static):
        raw = synth_cifar(n, seed) if cnv else synth_mnist(n, seed)
        d_raw = torch.from_numpy(raw).to(dev)
        d_in = torch.empty(n * rt.in_bytes, dtype=torch.uint8, device=dev)
        d_out = torch.zeros(n * rt.out_bytes, dtype=torch.uint8, device=dev)
        d_cls = torch.zeros(n, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        rt.prep_device(d_raw.data_ptr(), n, d_in.data_ptr())
        fn = rt.infer_device_static if static else rt.infer_device
        fn(d_in.data_ptr(), n, d_out.data_ptr(), d_cls.data_ptr() if cnv else 0, 10)
        rt.sync()
        return d_out.cpu().numpy(), d_cls.cpu().numpy()

    # before the device holds the parameters the entry point must refuse (the upload allocates; this call never does)
    with pytest.raises(RuntimeError, match="allocates nothing"):
        rt.infer_device_static(1, 1, 1)
    want1 = run(n1, 51, False)          # the existing entry point; also sets the device up
    want2 = run(n2, 52, False)
    before = rt.device_bytes(0)
    got1 = run(n1, 51, True)            # two calls in a row on the shared scratch, different inputs and sizes
    got2 = run(n2, 52, True)
    assert rt.device_bytes(0) == before, "the static entry point must not allocate device memory"
    for (wo, wc), (go, gc) in ((want1, got1), (want2, got2)):
        assert np.array_equal(go, wo)
        assert np.array_equal(gc, wc)
    assert np.array_equal(run(n1, 51, True)[0], want1[0])   # and once more, after a smaller call left its contents behind
    # beyond the stated maximum: refused with the limit in the message, nothing launched
    with pytest.raises(RuntimeError, match="exceed the static scratch.*at most %d images" % LIMIT[net]):
        rt.infer_device_static(1, LIMIT[net] + 1, 1)
    rt.close()
