#!/usr/bin/env python
"""on the GPU box, under compute-sanitizer: a few small host-buffer calls per network (every kernel family, ragged counts, the
CTA-pair instantiations) plus a short faulted call.  tools/sanitize_small.py [nets...]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle_py import params_dir, synth_cifar, synth_mnist
from bnn import native
for net in (sys.argv[1:] or ["cnvW1A1", "cnvW1A2", "lfcW1A1", "lfcW1A2"]):
    rt = native.Runtime(net, params_dir(net), devices=[0])
    cnv = net.startswith("cnv")
    for n in (1, 31, 300):
        data = synth_cifar(n, n) if cnv else synth_mnist(n, n)
        cls = rt.classify_cifar(data, 10) if cnv else rt.classify_mnist(data, 10)
        print(net, n, int(np.asarray(cls).sum()))
    fl = rt.schedule_faults(300, 8, word_size=1, target=-1, seed=3)
    data = synth_cifar(300, 9) if cnv else synth_mnist(300, 9)
    print(net, "faulted", int(rt.classify_with_fault_list(data, fl).sum()))
    rt.close()
print("sanitize_small: done")
