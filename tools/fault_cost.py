#!/usr/bin/env python
"""on the GPU box: cost of one fault epoch per target layer and memory (weights / thresholds), host-buffer path.
   tools/fault_cost.py [images] [faults]  -> microseconds added per fault = (t_faulted - t_clean) / faults"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
sys.path.insert(0, ROOT)
from bench import params_dir  # noqa: E402
from bnn import native  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 256
net = "cnvW1A1"
rt = native.Runtime(net, params_dir(net), devices=[0])
L = native.lib()
host, hp = native.pinned_array(n * 3073)
host[:] = np.random.default_rng(1).integers(0, 256, n * 3073, dtype=np.uint8)
cls, cp = native.pinned_array(n * 4)
us = C.c_float(0)


def call(fl):
    fp = fl.ctypes.data_as(C.c_void_p) if fl.shape[0] else None
    t0 = time.perf_counter()
    assert L.bnnb_classify_with_fault_list(rt.h, hp, n, 10, fp, fl.shape[0], cp, C.byref(us)) == 0, L.bnnb_last_error(rt.h)
    return time.perf_counter() - t0


empty = np.zeros((0, 5), np.uint32)
call(empty)
clean = min(call(empty) for _ in range(3))
print("clean: %.2f ms (%.2f M images/s)" % (clean * 1e3, n / clean / 1e6))
for target, name in ((0, "weights"), (1, "thresholds")):
    for layer in range(9):
        fl = rt.schedule_faults(n, nf, word_size=1, target=target, target_layers=[layer], seed=7)
        if fl.shape[0] == 0:
            continue
        rt.load_parameters(params_dir(net))
        call(empty)
        t = call(fl)
        print("%-10s layer %d: %4d faults  %.2f ms  -> %.1f us per fault" % (name, layer, fl.shape[0], t * 1e3, (t - clean) * 1e6 / fl.shape[0]))

# the bench's case: every layer and both memories in one schedule (BASELINE configs[4] density: one fault per 1 024 images)
for rep in range(3):
    fl = rt.schedule_faults(n, nf, word_size=1, target=-1, seed=3)
    rt.load_parameters(params_dir(net))
    call(empty)
    t = call(fl)
    grp = np.bincount(fl[:, 2].astype(np.int64), minlength=9)
    print("any layer (rep %d): %4d faults  %.2f ms  -> %.1f us per fault   faults per layer %s" % (rep, fl.shape[0], t * 1e3, (t - clean) * 1e6 / fl.shape[0], grp.tolist()))
