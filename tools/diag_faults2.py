#!/usr/bin/env python
"""on a 2-GPU box: where do the one-device and the all-devices faulted runs differ? (diagnostic)"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_py import params_dir, synth_cifar
from bnn import native
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
uniq = synth_cifar(2048, 105)
idx = np.random.default_rng(5).integers(0, 2048, n)
batch = np.ascontiguousarray(uniq.reshape(2048, 3073)[idx]).reshape(-1)
rt1 = native.Runtime("cnvW1A1", params_dir("cnvW1A1"), devices=[0])
rtg = native.Runtime("cnvW1A1", params_dir("cnvW1A1"))
rt1b = native.Runtime("cnvW1A1", params_dir("cnvW1A1"), devices=[0])
if os.environ.get("DIAG_UNFAULTED_FIRST"):
    u1 = rt1.classify_cifar(batch, 10); ug = rtg.classify_cifar(batch, 10)
    print("unfaulted agree:", np.array_equal(u1, ug))
fl = rt1.schedule_faults(n, 1024, word_size=1, target=-1, seed=11)
f1 = rt1.classify_with_fault_list(batch, fl)
f1b = rt1b.classify_with_fault_list(batch, fl)
fg = rtg.classify_with_fault_list(batch, fl)
print("one-device runs agree:", np.array_equal(f1, f1b))
d = np.nonzero(f1 != fg)[0]
print("differ:", d.size, "first", d[:10], "last", d[-5:] if d.size else None)
if d.size:
    t = np.sort(fl[:, 0])
    print("fault times around first diff:", t[np.searchsorted(t, d[0]) - 2: np.searchsorted(t, d[0]) + 2])
    print("layers/targets of faults just before:", fl[np.argsort(fl[:, 0])][np.searchsorted(t, d[0]) - 3: np.searchsorted(t, d[0]) + 1])
    h = np.histogram(d, bins=16, range=(0, n))[0]
    print("histogram of differing images over 16 bins:", h)
