"""CTA pairs (tcgen05 cta_group::2, conv_f4.cu): the paired and the single-CTA instantiations of every conv layer must give the
same scores, at image counts that leave the follower CTA of the last cluster without an image group, with one group only (no
pair launched), and at counts that are ragged against every layer's group size (9 / 6 / 15 images).  BNN_B200_CONV_PAIR is read
once per process, so each setting runs in a process of its own; the in-process result (the default mask) is checked against the
oracle as well."""
import hashlib
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle_py import NETS, Oracle, params_dir, synth_cifar

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))

pytestmark = pytest.mark.gpu
CNV = [n for n in NETS if n.startswith("cnv")]
SIZES = [1, 9, 10, 15, 16, 19, 31, 91, 277, 1000]  # 10 = one full group of layer 2 + 1 image; 16 = layer 4's group + 1; ...

CHILD = r"""
import hashlib, sys
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
from oracle_py import params_dir, synth_cifar
from bnn import native
net = sys.argv[3]
rt = native.Runtime(net, params_dir(net), devices=[0])
for n in [int(x) for x in sys.argv[4].split(",")]:
    cls, sc = rt.classify_cifar(synth_cifar(n, 100 + n), 10, want_scores=True)
    print(n, hashlib.sha256(np.ascontiguousarray(sc).tobytes()).hexdigest())
"""


def scores_by_mask(net, mask):
    env = dict(os.environ, BNN_B200_CONV_PAIR=str(mask), BNN_B200_QUIET="1")
    out = subprocess.run([sys.executable, "-c", CHILD, os.path.join(ROOT, "bnn-pynq_b200"), os.path.join(ROOT, "tests"), net,
                          ",".join(str(s) for s in SIZES)], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return dict(l.split() for l in out.stdout.strip().splitlines() if l and l[0].isdigit())


@pytest.mark.parametrize("net", CNV)
def test_paired_and_single_cta_conv_layers_agree(net):
    from bnn import native
    single, paired = scores_by_mask(net, 0), scores_by_mask(net, 7)
    assert single == paired and len(single) == len(SIZES)
    rt, o = native.Runtime(net, params_dir(net), devices=[0]), Oracle(net)
    for n in SIZES[:7]:  # the default mask, against the oracle
        rec = synth_cifar(n, 100 + n)
        cls, sc = rt.classify_cifar(rec, 10, want_scores=True)
        assert np.array_equal(sc, o.cnv_scores(rec)), n
        assert hashlib.sha256(np.ascontiguousarray(sc).tobytes()).hexdigest() == single[str(n)]
    rt.close()
