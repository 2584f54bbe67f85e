#!/usr/bin/env python
"""Condense ncu exports into the tracked summaries under profiles/.
  tools/ncu_summary.py launches <launches.csv> <out.csv>     per-kernel launch count, total device time, share of the run
  tools/ncu_summary.py full <report.ncu-rep> <out.csv>       one row per profiled launch with the metrics DESIGN.md quotes
  tools/ncu_summary.py traffic <full-summary.csv> <net> <images_per_launch>
                                                             DRAM bytes of the network's dominant kernel (longest launch of the
                                                             summary) -> profiles/traffic.json, which bench.py's roofline.traffic reads
"""
import json
import os
import csv
import subprocess
import sys
from collections import OrderedDict

FULL = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "gpu__time_duration.sum", "sm__cycles_elapsed.max",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active"]


def launches(src, dst):
    rows = [r for r in csv.reader(l for l in open(src) if not l.startswith("=="))]
    h = rows[0]
    ik, im, iv = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
    agg = OrderedDict()
    for r in rows[1:]:
        if len(r) <= iv or r[im] != "gpu__time_duration.sum":
            continue
        a = agg.setdefault(r[ik][:70], [0, 0.0])
        a[0] += 1
        a[1] += float(r[iv].replace(",", ""))
    # the path's own kernels (bench.py also runs pipe micro-benchmarks, an L2 flush and the load-time re-pack kernels)
    path = lambda k: any(x in k for x in ("k_cnv01", "k_conv", "k_gemm", "k_fc_chain", "k_layer_popc", "k_prep_", "k_cnv_classes"))
    tot = sum(a[1] for k, a in agg.items() if path(k))
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel", "launches", "total_ns", "share_of_path_kernels"])
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            w.writerow([k, a[0], int(a[1]), "%.4f" % (a[1] / tot) if path(k) else "-"])


def full(rep, dst):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h, u = rows[0], rows[1]
    cols = [h.index("Kernel Name")] + [h.index(m) for m in FULL if m in h]
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([h[c] for c in cols])
        w.writerow([u[c] for c in cols])
        for r in rows[2:]:
            w.writerow([r[c][:90] if c == cols[0] else r[c] for c in cols])


def traffic(summary_csv, net, images_per_launch):
    rows = list(csv.reader(open(summary_csv)))
    h, units = rows[0], rows[1]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1, "nsecond": 1, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "s": 1e9, "second": 1e9}
    num = lambda r, c: float(r[h.index(c)].replace(",", "")) * scale.get(units[h.index(c)], 1)   # the summary keeps ncu's units row
    body = [r for r in rows[2:] if len(r) == len(h)]
    top = max(body, key=lambda r: num(r, "gpu__time_duration.sum"))
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
    try:
        d = json.load(open(path))
    except Exception:
        d = {}
    d[net] = {"kernel": top[0], "dram_bytes_read": int(num(top, "dram__bytes_read.sum")), "dram_bytes_write": int(num(top, "dram__bytes_write.sum")),
              "images_per_launch": int(images_per_launch), "duration_ns": int(num(top, "gpu__time_duration.sum")),
              "source": "profiles/" + os.path.basename(summary_csv)}
    json.dump(d, open(path, "w"), indent=1, sort_keys=True)
    print(json.dumps(d[net]))


if __name__ == "__main__":
    if sys.argv[1] == "traffic":
        traffic(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
