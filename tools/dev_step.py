#!/usr/bin/env python
"""on the GPU box: N device-resident steps of B images and nothing else (for ncu captures whose totals must be divisible by an
exact image count).  tools/dev_step.py [net] [B] [steps]   (chunk size: $BNN_B200_CHUNK)"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))
sys.path.insert(0, ROOT)
from bench import params_dir  # noqa: E402
from bnn import native  # noqa: E402

net = sys.argv[1] if len(sys.argv) > 1 else "cnvW1A1"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
rt = native.Runtime(net, params_dir(net), devices=[0])
cnv = net.startswith("cnv")
raw_w = 3073 if cnv else 784
dev = torch.device("cuda", 0)
d_raw = torch.from_numpy(np.random.default_rng(1).integers(0, 256, B * raw_w, dtype=np.uint8)).to(dev)
d_in = torch.empty(B * rt.in_bytes, dtype=torch.uint8, device=dev)
d_out = torch.empty(B * rt.out_bytes, dtype=torch.uint8, device=dev)
d_cls = torch.empty(B, dtype=torch.int32, device=dev)
torch.cuda.synchronize()
rt.prep_device(d_raw.data_ptr(), B, d_in.data_ptr())
prof = os.environ.get("DEV_STEP_PROFILE")
if prof:  # per-layer CUDA-event times (after one warm-up step)
    rt.infer_device(d_in.data_ptr(), B, d_out.data_ptr(), d_cls.data_ptr() if cnv else 0, 10)
    rt.sync()
    rt.profile_enable(True)
for _ in range(steps):
    rt.infer_device(d_in.data_ptr(), B, d_out.data_ptr(), d_cls.data_ptr() if cnv else 0, 10)
rt.sync()
if prof:
    ms, cnt = rt.profile_collect()
    print(prof, net, "layer_ms", [round(float(x) / steps, 3) for x in ms], "total %.3f" % (float(sum(ms)) / steps))
print("done: %d steps x %d images" % (steps, B))
