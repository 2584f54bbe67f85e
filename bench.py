#!/usr/bin/env python
"""bench.py -- CNV-W1A1 images/s on B200 (BASELINE.json metric), one process per GPU.

A "step" is one pass of the hot path (BlackBoxJam compute mode: layer-0 input -> 64 x int16 scores + class) over one batch
of 65 536 synthetic CIFAR-10-shaped images per GPU with the shipped cifar10/cnvW1A1 parameters.
  value : images/s with the int8 HWC inputs already resident in HBM (the reference's own timed region,
          foldedmv-offload.h:389-392), CUDA events on the launching stream, max over ranks.
  e2e   : the same metric through the C ABI with HOST buffers (bnnb_classify_cifar): pinned records -> H2D -> on-device
          quantise/interleave -> layers -> classes D2H, wall clock around the synchronous call.
  roofline / cpu_baseline / clocks / gpu_launches: see DESIGN.md "Measurement".
`--impl reference` times the reference's own sw runtime (oracle/_ref, built from the unmodified sources) on the host cores.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200"))

ALG_BYTES = {"cnv": 3072 + 128, "lfc": 104 + 8}          # what BlackBoxJam reads + writes per image
OPS = {"cnv": 118978048, "lfc": 6029312}                  # config.h #Ops per image (2 x MAC)
PARAMS = {"cnvW1A1": "cifar10/cnvW1A1", "cnvW1A2": "cifar10/cnvW1A2", "cnvW2A2": "cifar10/cnvW2A2",
          "lfcW1A1": "mnist/lfcW1A1", "lfcW1A2": "mnist/lfcW1A2"}


def params_dir(net):
    return os.path.join(ROOT, "tests", "golden", "params", PARAMS[net])


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(s) > 2 + i and s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}
