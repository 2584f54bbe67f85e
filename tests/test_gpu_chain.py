"""The trailing fully-connected layers in one launch (fc_chain.cu: CNV layers 6-8, LFC layers 1-3) against the same layers launched
one by one (gemm_f4.cu, BNN_B200_CHAIN=0), for all five networks: scores / output words and classes must be identical at image
counts around the 128-image tile (1, 127, 128, 129), with several tiles per CTA (20 000 images on 148 SMs = 157 tiles) and at a
ragged multi-tile count.  BNN_B200_CHAIN is read once per engine, so each setting runs in a process of its own; the in-process
result (chain on, the default) is checked against the oracle as well."""
import hashlib
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle_py import NETS, Oracle, params_dir, synth_cifar, synth_mnist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))

pytestmark = pytest.mark.gpu
SIZES = [1, 127, 128, 129, 1000, 20000]

CHILD = r"""
import hashlib, sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
from oracle_py import params_dir, synth_cifar, synth_mnist
from bnn import native
net = sys.argv[3]
rt = native.Runtime(net, params_dir(net), devices=[0])
for n in [int(x) for x in sys.argv[4].split(",")]:
    if net.startswith("cnv"):
        cls, raw = rt.classify_cifar(synth_cifar(n, 300 + n), 10, want_scores=True)
    else:
        cls, raw = rt.classify_mnist(synth_mnist(n, 300 + n), 10, want_words=True)
    h = hashlib.sha256(np.ascontiguousarray(raw).tobytes() + np.ascontiguousarray(cls).tobytes()).hexdigest()
    print(n, h)
"""


def results(net, chain):
    env = dict(os.environ, BNN_B200_CHAIN=str(chain), BNN_B200_QUIET="1")
    out = subprocess.run([sys.executable, "-c", CHILD, os.path.join(ROOT, "bnn-pynq_b200"), os.path.join(ROOT, "tests"), net,
                          ",".join(str(s) for s in SIZES)], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return dict(l.split() for l in out.stdout.strip().splitlines() if l and l[0].isdigit())


@pytest.mark.parametrize("net", NETS)
def test_chained_and_separate_fc_layers_agree(net):
    from bnn import native
    one_by_one, chained = results(net, 0), results(net, 1)
    assert one_by_one == chained and len(chained) == len(SIZES)
    rt, o = native.Runtime(net, params_dir(net), devices=[0]), Oracle(net)
    for n in SIZES[:5]:  # the default (chain on), against the oracle
        if net.startswith("cnv"):
            rec = synth_cifar(n, 300 + n)
            cls, raw = rt.classify_cifar(rec, 10, want_scores=True)
            want = o.cnv_scores(rec)
        else:
            px = synth_mnist(n, 300 + n)
            cls, raw = rt.classify_mnist(px, 10, want_words=True)
            want = o.lfc_words(px)
        assert np.array_equal(raw, want), n
        h = hashlib.sha256(np.ascontiguousarray(raw).tobytes() + np.ascontiguousarray(cls).tobytes()).hexdigest()
        assert h == chained[str(n)]
    rt.close()


def test_chain_is_the_path_that_runs():
    """The launch count of a device-resident pass: CNV = fused 0+1, layers 2, 3, 4, 5 and the chain (6 kernels); with
    BNN_B200_CHAIN=0 two more.  Guards against the chain silently refusing its launch."""
    code = ("import sys; sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2]);"
            "import numpy as np, torch;"
            "from oracle_py import params_dir, synth_cifar; from bnn import native;"
            "rt = native.Runtime('cnvW1A1', params_dir('cnvW1A1'), devices=[0]); B = 256;"
            "raw = torch.from_numpy(np.ascontiguousarray(synth_cifar(B, 1)).reshape(-1)).cuda();"
            "d_in = torch.empty(B * rt.in_bytes, dtype=torch.uint8, device='cuda');"
            "d_out = torch.empty(B * rt.out_bytes, dtype=torch.uint8, device='cuda');"
            "d_cls = torch.empty(B, dtype=torch.int32, device='cuda'); torch.cuda.synchronize();"
            "rt.prep_device(raw.data_ptr(), B, d_in.data_ptr());"
            "rt.infer_device(d_in.data_ptr(), B, d_out.data_ptr(), d_cls.data_ptr(), 10); rt.sync();"
            "print('launches', rt.last_launches)")
    counts = {}
    for chain in (0, 1):
        env = dict(os.environ, BNN_B200_CHAIN=str(chain), BNN_B200_QUIET="1")
        out = subprocess.run([sys.executable, "-c", code, os.path.join(ROOT, "bnn-pynq_b200"), os.path.join(ROOT, "tests")], env=env,
                             capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr[-2000:]
        counts[chain] = int(out.stdout.split("launches")[1].split()[0])
    assert counts[0] - counts[1] == 2, counts
