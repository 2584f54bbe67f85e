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
LAYER_OPS = {"cnv": [3110400, 57802752, 21233664, 29491200, 5308416, 1179648, 262144, 524288, 65536],   # cnvW1A1/hw/config.h:15..190
             "lfc": [1703936, 2097152, 2097152, 131072]}                                                  # lfcW1A1/hw/config.h:14..71
PARAMS = {"cnvW1A1": "cifar10/cnvW1A1", "cnvW1A2": "cifar10/cnvW1A2", "cnvW2A2": "cifar10/cnvW2A2",
          "lfcW1A1": "mnist/lfcW1A1", "lfcW1A2": "mnist/lfcW1A2"}


def params_dir(net):
    return os.path.join(ROOT, "tests", "golden", "params", PARAMS[net])


def peaks():
    """(hbm GB/s, dense bf16 TFLOP/s burst, source).  int8 tensor peak is nominally 2x the bf16 figure."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
            return float(d["hbm_gbs"]), float(d["bf16_tflops"]), "measured"
    except Exception:
        return 6650.0, 1590.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], False

    def _nvml(self):
        """Fast path: NVML in-process (a few samples per millisecond-scale step instead of one nvidia-smi fork per 100 ms)."""
        import pynvml as N
        N.nvmlInit()
        h = N.nvmlDeviceGetHandleByIndex(self.index)
        mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
        bits = [(N.nvmlClocksEventReasonHwSlowdown, 2), (N.nvmlClocksEventReasonHwThermalSlowdown, 3),
                (N.nvmlClocksEventReasonSwThermalSlowdown, 4), (N.nvmlClocksEventReasonSwPowerCap, 5)]
        while not self.stop_flag:
            sm = N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)
            r = N.nvmlDeviceGetCurrentClocksEventReasons(h)
            row = [str(sm), str(mx), "", "", "", ""]
            for bit, pos in bits:
                row[pos] = "Active" if (r & bit) else "Not Active"
            self.samples.append(row)
            time.sleep(0.002)

    def run(self):
        try:
            self._nvml()
            return
        except Exception:
            pass
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


def synth_records(n, seed):
    import numpy as np
    rng = np.random.default_rng(seed)
    rec = np.zeros((n, 3073), np.uint8)
    rec[:, 1:] = rng.integers(0, 256, (n, 3072), dtype=np.uint8)
    return rec


# ---------------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own sw runtime (oracle/_ref/python_sw-<net>.so), one process per host core
# ---------------------------------------------------------------------------------------------------------------
def _ref_worker(args):
    net, pdir, nimg, seed, steps = args
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)  # the reference prints per call
    from ref_py import Ref
    r = Ref(net, pdir)
    rec = synth_records(nimg, seed).tobytes()
    n = C.c_int(0)
    us = C.c_float(0)
    path = r._file(rec)
    out = []
    for _ in range(steps):
        t0 = time.perf_counter()
        p = r.L.inference_multiple(path.encode(), 10, C.byref(n), C.byref(us), 0)
        wall = time.perf_counter() - t0
        r.L.free_results(p)
        out.append((us.value * n.value * 1e-6, wall))  # the reference's own timed region (BlackBoxJam only), and wall
    os.unlink(path)
    return out


def run_reference_cpu(net, per_proc, steps, warmup, procs=None):
    """Returns (images/s over the timed steps, ms per step, cores, sample description)."""
    from multiprocessing import get_context
    so = os.path.join(ROOT, "oracle", "_ref", "python_sw-%s.so" % net)
    if not os.path.exists(so):
        raise RuntimeError("oracle/_ref is missing (built by oracle/build_ref.sh where /root/reference exists)")
    procs = procs or len(os.sched_getaffinity(0))
    with get_context("spawn").Pool(procs) as pool:
        res = pool.map(_ref_worker, [(net, params_dir(net), per_proc, 100 + i, steps + warmup) for i in range(procs)])
    # per step: all processes run concurrently; the step lasts as long as the slowest accelerator call
    step_s = [max(res[p][k][0] for p in range(procs)) for k in range(warmup, warmup + steps)]
    total = sum(step_s)
    ips = procs * per_proc * steps / total
    sample = "%d images per process x %d processes per step, reference BlackBoxJam region (foldedmv-offload.h:389-392)" % (per_proc, procs)
    return ips, 1000.0 * total / steps, procs, sample


def _peaks_json():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def one_process_e2e(native, L, net, ngpu, B, steps):
    """Host-buffer images/s of ONE process sharding ngpu x B images over ngpu devices through bnnb_classify_cifar (Engine::infer_host:
    contiguous shards, one feeder thread per device)."""
    import numpy as np
    rt = native.Runtime(net, params_dir(net), devices=list(range(ngpu)))
    n = ngpu * B
    host, host_ptr = native.pinned_array(n * 3073)
    rng = np.random.default_rng(9)
    for o in range(0, n, 65536):
        m = min(65536, n - o)
        host[o * 3073:(o + m) * 3073] = rng.integers(0, 256, m * 3073, dtype=np.uint8)
    cls, cls_ptr = native.pinned_array(n * 4)
    us = C.c_float(0)

    def call():
        if L.bnnb_classify_cifar(rt.h, host_ptr, n, 10, cls_ptr, None, C.byref(us)) != 0:
            raise RuntimeError(L.bnnb_last_error(rt.h).decode())

    call(); call()
    t0 = time.perf_counter()
    for _ in range(steps):
        call()
    dt = (time.perf_counter() - t0) / steps
    native.free_pinned(host_ptr); native.free_pinned(cls_ptr)
    rt.close()
    return {"value": n / dt, "unit": "images/s", "images_per_call": n, "devices": ngpu, "device_kernel_usec_per_image": float(us.value)}


def traffic_per_image(net):
    """dram__bytes_read.sum + dram__bytes_write.sum of the network's dominant kernel PER IMAGE, from profiles/traffic.json, which
    tools/ncu_summary.py writes from an `ncu --set full` capture (one entry per network: bytes of ONE launch and the images that launch
    processed).  None where no capture exists.  roofline.traffic = this x the images one launch processes."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            e = json.load(f)[net]
        return (e["dram_bytes_read"] + e["dram_bytes_write"]) / float(e["images_per_launch"]), e.get("source")
    except Exception:
        return None, None


def fast_cpu_port(net, seconds=10.0):
    """Second CPU figure: the multi-threaded C restatement oracle/libbnn_oracle.so - NOT reference code - on all host
    cores, on a bounded sample of the workload.  Part of the cpu_baseline leg (the only place bench.py may execute oracle/)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_py import Oracle
    import numpy as np
    o = Oracle(net, params_dir(net))
    cnv = net.startswith("cnv")
    n = 256 if cnv else 8192
    rng = np.random.default_rng(7)
    if cnv:
        data = np.zeros((n, 3073), np.uint8)
        data[:, 1:] = rng.integers(0, 256, (n, 3072), dtype=np.uint8)
        run = o.cnv_scores
    else:
        data = rng.integers(0, 256, (n, 784), dtype=np.uint8)
        run = o.lfc_words
    run(data[:16])
    t0 = time.perf_counter()
    run(data)
    dt = time.perf_counter() - t0
    reps = max(1, min(64, int(seconds / max(dt, 1e-6))))
    t0 = time.perf_counter()
    for _ in range(reps):
        run(data)
    dt = time.perf_counter() - t0
    cores = len(os.sched_getaffinity(0))
    return {"value": reps * n / dt, "unit": "images/s", "cores": cores, "kind": "port",
            "sample": "%d images x %d passes, OpenMP over %d cores" % (n, reps, cores),
            "note": "oracle/bnn_oracle.c, the int8 restatement used as the parity checker: not reference code"}


def measure_sustained(step, seconds, stream_events, B, world, max_over_ranks, barrier, sampler_cls, local_rank):
    """Back-to-back steps for `seconds` (no L2 flush, no host work between launches): the sustained arm, with its own clock samples."""
    import torch
    probe_a, probe_b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    probe_a.record(); step(); probe_b.record(); torch.cuda.synchronize()
    per = max(probe_a.elapsed_time(probe_b), 1e-3)
    steps = max(10, int(seconds * 1000.0 / per))
    sampler = sampler_cls(local_rank)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    sampler.stop_flag = True
    sampler.join()
    ms = max_over_ranks(e0.elapsed_time(e1))
    return {"seconds": ms * 1e-3, "steps": steps, "ms_per_step": ms / steps, "value": world * B * steps / (ms * 1e-3), "unit": "images/s",
            "clocks": sampler.summary()}


def measure_latency(L, rt, fam, host_ptr, calls):
    """Batch-1 latency of the host-buffer call bnn.py's inference() makes (record in pinned host memory -> class on the host)."""
    import numpy as np
    from bnn import native
    out, out_ptr = native.pinned_array(64)
    us = C.c_float(0)
    fn = L.bnnb_classify_cifar if fam == "cnv" else L.bnnb_classify_mnist
    ts = []
    for i in range(calls + 20):
        t0 = time.perf_counter()
        rc = fn(rt.h, host_ptr, 1, 10, out_ptr, None, C.byref(us))
        ts.append(time.perf_counter() - t0)
        if rc != 0:
            raise RuntimeError(L.bnnb_last_error(rt.h).decode())
    ts = np.sort(np.array(ts[20:])) * 1e6
    native.free_pinned(out_ptr)
    return {"calls": calls, "median_us": float(ts[len(ts) // 2]), "p10_us": float(ts[len(ts) // 10]), "p99_us": float(ts[int(len(ts) * 0.99)]),
            "device_kernel_us": float(us.value),
            "reference_fpga_us": {"cnvW1A1": 527, "cnvW1A2": 543, "cnvW2A2": 1622, "lfcW1A1": 8}.get(rt.network),
            "note": "wall clock around bnnb_classify_* with n = 1 (H2D of one record, quantise, all layers, class D2H); the FPGA figure is the "
                    "reference's notebook number for its overlay (notebooks/CNV-QNN_Cifar10.ipynb:129,217,298, LFC-BNN_MNIST_Webcam.ipynb:291)"}


def measure_faults(native, L, net, local_rank, rank, world, total_images, nfaults, barrier, max_over_ranks):
    """BASELINE configs[4]: `total_images` synthetic images, sharded contiguously over the ranks, classified through
    bnnb_classify_with_fault_list with a seeded schedule of `nfaults` parameter upsets (time i = applied before image i of the WHOLE job:
    a rank applies the faults that fall before its shard first, then runs its segments) - against the same call with no faults."""
    import numpy as np
    fam_cnv = net.startswith("cnv")
    raw_w = 3073 if fam_cnv else 784
    lo, hi = rank * total_images // world, (rank + 1) * total_images // world
    n = hi - lo
    rt = native.Runtime(net, params_dir(net), devices=[local_rank])
    host, host_ptr = native.pinned_array(n * raw_w)
    rng = np.random.default_rng(2 + rank)
    blk = 1 << 16
    for o in range(0, n, blk):   # generated in blocks: 3 GB of random bytes at once would double the host footprint
        m = min(blk, n - o)
        r = rng.integers(0, 256, (m, raw_w), dtype=np.uint8)
        if fam_cnv:
            r[:, 0] = 0
        host[o * raw_w:(o + m) * raw_w] = r.reshape(-1)
    cls, cls_ptr = native.pinned_array(n * 4)
    fl = rt.schedule_faults(total_images, nfaults, word_size=1, target=-1, seed=3)   # the same list on every rank
    mine = fl.copy()
    t = mine[:, 0].astype(np.int64) - lo
    mine[:, 0] = np.clip(t, 0, max(n - 1, 0)).astype(np.uint32)   # earlier faults: before my first image
    late = t >= n                                                 # later faults: after my last image (they persist, nothing to classify)
    run_list = np.ascontiguousarray(mine[~late])
    us = C.c_float(0)

    def call(flist):
        fp = flist.ctypes.data_as(C.c_void_p) if flist.shape[0] else None
        rc = L.bnnb_classify_with_fault_list(rt.h, host_ptr, n, 10, fp, flist.shape[0], cls_ptr, C.byref(us))
        if rc != 0:
            raise RuntimeError(L.bnnb_last_error(rt.h).decode())

    empty = np.zeros((0, 5), np.uint32)
    call(empty)                      # warm-up (buffers, clocks)
    barrier()
    t0 = time.perf_counter(); call(empty); t_clean = max_over_ranks(time.perf_counter() - t0)
    clean = cls.view(np.int32).copy()
    rt.load_parameters(params_dir(net))
    barrier()
    t0 = time.perf_counter(); call(run_list); t_cold = max_over_ranks(time.perf_counter() - t0)   # first faulted call: parameter clones, kernel loading
    rt.load_parameters(params_dir(net))   # discards the faults: the timed call starts from the shipped parameters again
    barrier()
    t0 = time.perf_counter(); call(run_list); t_fault = max_over_ranks(time.perf_counter() - t0)
    faulted = cls.view(np.int32).copy()
    first = int(run_list[:, 0].min()) if run_list.shape[0] else n
    ok_prefix = bool(np.array_equal(clean[:first], faulted[:first]))
    changed = int((clean != faulted).sum())
    segs = int(np.unique(run_list[:, 0]).size) + 1 if run_list.shape[0] else 1
    native.free_pinned(host_ptr); native.free_pinned(cls_ptr)
    rt.close()
    return {"images": total_images, "faults_scheduled": int(fl.shape[0]), "segments_rank0": segs,
            "e2e_unfaulted_images_per_s": total_images / t_clean, "e2e_faulted_images_per_s": total_images / t_fault,
            "e2e_faulted_first_call_images_per_s": total_images / t_cold,
            "ratio": t_clean / t_fault, "results_changed_rank0": changed, "prefix_before_first_fault_identical": ok_prefix,
            "note": "bnnb_classify_with_fault_list on pinned host records, wall clock, max over ranks; bit-exactness of faulted runs is pinned by "
                    "tests/test_gpu_full_size.py (oracle) - bench.py only checks that nothing changes before the first fault time"}


def measure_network(native, L, net, local_rank, steps, B):
    """One more network, briefly: device-resident images/s and host-buffer images/s (side field `networks`)."""
    import numpy as np
    import torch
    rt = native.Runtime(net, params_dir(net), devices=[local_rank])
    cnv = net.startswith("cnv")
    raw_w = 3073 if cnv else 784
    host, host_ptr = native.pinned_array(B * raw_w)
    host[:] = np.random.default_rng(5).integers(0, 256, B * raw_w, dtype=np.uint8)
    dev = torch.device("cuda", local_rank)
    d_raw = torch.from_numpy(host).to(dev)
    d_in = torch.empty(B * rt.in_bytes, dtype=torch.uint8, device=dev)
    d_out = torch.empty(B * rt.out_bytes, dtype=torch.uint8, device=dev)
    d_cls = torch.empty(B, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    rt.prep_device(d_raw.data_ptr(), B, d_in.data_ptr(), stream)

    def step():
        rt.infer_device(d_in.data_ptr(), B, d_out.data_ptr(), d_cls.data_ptr() if cnv else 0, 10, stream)

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    rt.sync()
    ms = e0.elapsed_time(e1) / steps
    cls_host, cls_ptr = native.pinned_array(B * 4)
    words_host, words_ptr = native.pinned_array(B * 8)
    us = C.c_float(0)

    def e2e():
        rc = (L.bnnb_classify_cifar(rt.h, host_ptr, B, 10, cls_ptr, None, C.byref(us)) if cnv
              else L.bnnb_classify_mnist(rt.h, host_ptr, B, 10, cls_ptr, words_ptr, C.byref(us)))
        if rc != 0:
            raise RuntimeError(L.bnnb_last_error(rt.h).decode())

    e2e()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e()
    dt = (time.perf_counter() - t0) / steps
    fam = "cnv" if cnv else "lfc"
    res = {"value": B / (ms * 1e-3), "ms_per_step": ms, "batch": B, "e2e": B / dt, "unit": "images/s", "tops": OPS[fam] * B / (ms * 1e-3) / 1e12}
    if not cnv:  # the reference packs MNIST on the host (binarizeAndPack): the same call on 104-byte packed inputs
        pk, pk_ptr = native.pinned_array(B * 104)
        pk[:] = native.pack_mnist(host.reshape(B, 784)).reshape(-1)

        def e2e_packed():
            if L.bnnb_classify_packed(rt.h, pk_ptr, B, 10, cls_ptr, None, words_ptr, C.byref(us)) != 0:
                raise RuntimeError(L.bnnb_last_error(rt.h).decode())

        e2e_packed()
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_packed()
        res["e2e_packed_input"] = B / ((time.perf_counter() - t0) / steps)
        res["e2e_packed_input_note"] = "bnnb_classify_packed: 104 B per image over PCIe (host-side binarizeAndPack, as the reference does) instead of 784"
        native.free_pinned(pk_ptr)
    for ptr in (host_ptr, cls_ptr, words_ptr):
        native.free_pinned(ptr)
    rt.close()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--net", default="cnvW1A1")
    ap.add_argument("--batch", type=int, default=65536, help="images per GPU per step")
    ap.add_argument("--ref-images", type=int, default=8, help="reference arm: images per process per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--faults", type=int, default=1024, help="config 5 side measurement: faults scheduled over 1 048 576 images (0 = skip)")
    ap.add_argument("--fault-images", type=int, default=1 << 20, help="config 5 side measurement: images in the whole job (sharded over the ranks)")
    ap.add_argument("--sustained-s", type=float, default=2.5, help="sustained arm: seconds of back-to-back steps (0 = skip)")
    ap.add_argument("--latency-calls", type=int, default=1000, help="batch-1 latency: calls through the host-buffer C ABI (0 = skip)")
    ap.add_argument("--nets", default="all", help="comma list of further networks measured briefly into the `networks` side field ('all' = the other four)")
    ap.add_argument("--lean", action="store_true", help="only the headline measurement (kernel experiments): no side fields, no CPU baselines")
    ap.add_argument("--core", type=int, default=2, choices=[0, 1, 2],
                    help="compute core: 2 = tcgen05 kind::mxf4 (default), 1 = tcgen05 kind::i8, 0 = LOP3+POPC bit-serial (BASELINE configs[3] comparison)")
    a = ap.parse_args()
    if a.lean:
        a.faults, a.sustained_s, a.latency_calls, a.nets, a.no_cpu_baseline = 0, 0.0, 0, "", True
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    fam = "cnv" if a.net.startswith("cnv") else "lfc"
    metric = "%s images/sec" % a.net.upper().replace("CNVW", "CNV-W").replace("LFCW", "LFC-W")
    config = {"workload": "%s, shipped %s params, batch %d synthetic images per GPU per step (uniform random bytes, seed 1)" % (a.net, PARAMS[a.net], a.batch),
              "batch_per_gpu": a.batch, "sharding": "contiguous image ranges, parameters replicated, no collective",
              "core": {2: "tcgen05 kind::mxf4", 1: "tcgen05 kind::i8", 0: "LOP3+POPC bit-serial"}[a.core]}
    in_bytes_step = a.batch * (3072 if fam == "cnv" else 104)
    flush_l2 = in_bytes_step <= 192 * 1024 * 1024   # inputs that could sit in the 126 MB L2: flush between timed steps
    config["l2"] = ("inputs (%d MB per step) could stay in the 126 MB L2: a 256 MB buffer is overwritten between timed steps (outside the events)"
                    if flush_l2 else "inputs (%d MB per step) exceed the 126 MB L2; no explicit flush") % (in_bytes_step >> 20)

    if a.impl == "reference":
        if rank != 0:
            return
        ips, ms, cores, sample = run_reference_cpu(a.net, a.ref_images, a.steps, max(a.warmup, 1))
        print(json.dumps({"impl": "reference", "metric": metric, "value": ips, "unit": "images/s", "n_gpus": a.gpus, "steps": a.steps,
                          "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "dtype": "int1/int16 (HLS C++ on CPU)", "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": "reference", "sample": sample},
                          "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from bnn import native
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the runtime has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's own banner ("NCCL version ...") goes to stderr.  NCCL initialises lazily,
        # so file descriptor 1 is pointed at stderr across the init AND the first collective.
        os.environ["NCCL_DEBUG"] = os.environ.get("BENCH_NCCL_DEBUG", "WARN")
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            t = torch.zeros(1, device=dev)
            dist.all_reduce(t)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    all_cpus = os.sched_getaffinity(0)
    numa_cpus = native.bind_host_to_device(local_rank) if not os.environ.get("BENCH_NO_BIND") else None  # before any pinned allocation
    rt = native.Runtime(a.net, params_dir(a.net), devices=[local_rank])
    if a.core != 2:
        rt.set_core(a.core)
    B = a.batch
    raw_w = 3073 if fam == "cnv" else 784
    host, host_ptr = native.pinned_array(B * raw_w)
    if fam == "cnv":
        host[:] = synth_records(B, 1 + rank).reshape(-1)
    else:
        host[:] = np.random.default_rng(1 + rank).integers(0, 256, B * 784, dtype=np.uint8)
    d_raw = torch.empty(B * raw_w, dtype=torch.uint8, device=dev)
    d_raw.copy_(torch.from_numpy(host), non_blocking=False)
    d_in = torch.empty(B * rt.in_bytes, dtype=torch.uint8, device=dev)
    d_out = torch.empty(B * rt.out_bytes, dtype=torch.uint8, device=dev)
    d_cls = torch.empty(B, dtype=torch.int32, device=dev)
    # a real (non-default) stream: the library launches on it and the CUDA events below are recorded on it
    tstream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(tstream)
    stream = tstream.cuda_stream
    assert stream != 0
    rt.prep_device(d_raw.data_ptr(), B, d_in.data_ptr(), stream)
    torch.cuda.synchronize()

    def step():
        rt.infer_device(d_in.data_ptr(), B, d_out.data_ptr(), d_cls.data_ptr() if fam == "cnv" else 0, 10, stream)

    for _ in range(max(a.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    rt.profile_enable(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if flush_l2:
        scratch = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
        for ea, eb in evs:
            scratch.fill_(1)          # same stream, before the start event
            ea.record()
            step()
            eb.record()
        barrier()
        ms_total = max_over_ranks(sum(ea.elapsed_time(eb) for ea, eb in evs))
    else:
        e0.record()
        for _ in range(a.steps):
            step()
        e1.record()
        barrier()
        ms_total = max_over_ranks(e0.elapsed_time(e1))
    rt.profile_enable(False)
    rt.sync()  # waits for the device and fails loudly if a kernel gave up waiting on an mbarrier during the timed region
    launches_per_step = rt.last_launches
    lay_ms, lay_cnt = rt.profile_collect()
    sampler.stop_flag = True
    sampler.join()
    ms_step = ms_total / a.steps
    value = world * B * a.steps / (ms_total * 1e-3)
    dev_classes = d_cls.cpu().numpy() if fam == "cnv" else None

    # ---- end to end through the C ABI with host buffers -------------------------------------------------------
    cls_host, cls_ptr = native.pinned_array(B * 4)
    words_host, words_ptr = native.pinned_array(B * 8)
    L = native.lib()
    us = C.c_float(0)

    def e2e_step():
        if fam == "cnv":
            rc = L.bnnb_classify_cifar(rt.h, host_ptr, B, 10, cls_ptr, None, C.byref(us))
        else:
            rc = L.bnnb_classify_mnist(rt.h, host_ptr, B, 10, cls_ptr, words_ptr, C.byref(us))
        if rc != 0:
            raise RuntimeError(L.bnnb_last_error(rt.h).decode())

    # plain pinned-host -> device copy of one step's input: the floor of the end-to-end number
    th = torch.from_numpy(host)
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    d_raw.copy_(th, non_blocking=True)
    torch.cuda.synchronize()
    h0.record()
    for _ in range(3):
        d_raw.copy_(th, non_blocking=True)
    h1.record()
    torch.cuda.synchronize()
    h2d_gbs = 3 * B * raw_w / (h0.elapsed_time(h1) * 1e-3) / 1e9

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_value = world * B * a.steps / e2e_s
    if fam == "cnv" and not os.environ.get("BNNB_BENCH_NOCHECK"):  # (kernel-stripping experiments produce garbage on purpose)
        assert np.array_equal(cls_host.view(np.int32), dev_classes), "host-buffer and device-resident paths disagree"

    # ---- roofline of the dominant kernel (live CUDA-event durations over the timed region) ----------------------
    # The path is compute-bound by three orders of magnitude (3 200 B and 119 M ops per image: HBM allows 2.05 G images/s,
    # the integer / tensor pipes a few M): the bound is the tensor pipe where a layer runs on it, and the same TOP/s scale is
    # used for CUDA-core (LOP3+POPC) layers so that fractions are comparable.  1 MAC = 2 ops, as config.h counts them.
    hbm_peak, tf_peak, how = peaks()
    lops = list(LAYER_OPS[fam])
    fused = []
    for i in range(len(lops) - 1):   # a layer without its own launch was fused into the next kernel: move its ops there
        if lay_cnt[i] == 0 and lops[i] and lay_cnt[i + 1] > 0:
            lops[i + 1] += lops[i]; lops[i] = 0; fused.append(i)
    k = int(np.argmax(lay_ms))
    launches_k = max(int(lay_cnt[k]), 1)
    units_per_launch = B * a.steps / launches_k
    avg_ms = float(lay_ms[k]) / launches_k
    achieved = lops[k] * units_per_launch / (avg_ms * 1e-3) / 1e12
    popc, _ = native.microbench(0)
    imma, _ = native.microbench(1)
    # kind::mxf4 issues at 4x the bf16 rate (nominal 9 vs 2.25 PFLOP/s dense): the tensor peak this path is held against is
    # 4 x the MEASURED dense bf16 number; the fraction of the plain bf16 number is kept beside it
    f4_peak = 4.0 * tf_peak
    tpi, tsrc = traffic_per_image(a.net)
    roofline = {"bound": "tensor", "achieved": achieved, "peak": f4_peak, "unit": "TFLOP/s", "frac": achieved / f4_peak, "traffic": (int(tpi * units_per_launch) if tpi else None), "traffic_source": tsrc,
                "kernel": "layer %d%s" % (k, " (fused with layer %d)" % (k - 1) if (k - 1) in fused else ""), "avg_launch_ms": avg_ms, "images_per_launch": units_per_launch,
                "ops_per_image_this_kernel": lops[k],
                "peak_source": "4 x " + how + " dense bf16 (%.1f TFLOP/s): tcgen05 kind::mxf4 nominal rate; ops = 2 x MAC as config.h counts them" % tf_peak,
                "frac_of_bf16_peak": achieved / tf_peak,
                "shape_ceiling": "layer 1 is M128.N64.K64 (64 output channels): SS-mode issue rate 6850 MAC/clk/SM = 3.98 POP/s per GPU (tests/test_gpu_tcgen05.py rate report)",
                "share_of_step": float(lay_ms[k] / max(lay_ms.sum(), 1e-9)),
                "layer_ms_per_step": [float(x) / a.steps for x in lay_ms],
                "layer_tops": [lops[i] * B * a.steps / (float(lay_ms[i]) * 1e-3) / 1e12 if lay_ms[i] > 0 else None for i in range(len(lops))],
                "step": {"tops": OPS[fam] * value / world / 1e12, "frac": OPS[fam] * value / world / 1e12 / f4_peak, "frac_of_bf16_peak": OPS[fam] * value / world / 1e12 / tf_peak,
                         "hbm_gbs": ALG_BYTES[fam] * value / world / 1e9, "hbm_frac": ALG_BYTES[fam] * value / world / 1e9 / hbm_peak},
                "pipes": {"popc_per_s_measured": popc, "popc_roofline_images_per_s": popc * 32 / (OPS[fam] / 2),
                          "popc_pipe_frac_step": (OPS[fam] / 2 / 32) * (value / world) / popc,
                          "int8_mma_sync_mac_per_s_measured": imma}}

    out = {"metric": metric, "value": value, "unit": "images/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
           "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "int1 x int1 as e2m1 on tcgen05 kind::mxf4 with unit scales (exact; fp32 accumulate), layer 0 int8 (int32 accumulate)" if a.net.endswith("W1A1") else "int2 (ternary activations, 1/2-bit weights) as e2m1 on tcgen05 kind::mxf4 with unit scales (exact; fp32 accumulate)", "data": "synthetic", "config": config,
           "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": B * raw_w, "d2h_bytes_per_step": B * (4 if fam == "cnv" else 8),
                   "h2d_copy_gbs_measured": h2d_gbs, "h2d_bound_images_per_s": world * h2d_gbs * 1e9 / raw_w,
                   "host_cpus_bound_to_gpu_node": (len(numa_cpus) if numa_cpus else None)},
           "gpu_launches": int(launches_per_step) * a.steps, "roofline": roofline, "clocks": sampler.summary()}
    # ---- side measurements (each optional; none of them changes the headline fields above) ---------------------------
    tf_sust = _peaks_json().get("bf16_tflops_sustained")
    if a.sustained_s > 0:
        sus = measure_sustained(step, a.sustained_s, None, B, world, max_over_ranks, barrier, ClockSampler, local_rank)
        rt.sync()
        sus["tops"] = OPS[fam] * sus["value"] / world / 1e12
        if tf_sust:
            sus["frac_of_4x_bf16_sustained"] = sus["tops"] / (4.0 * float(tf_sust))
            sus["peak"] = 4.0 * float(tf_sust)
        sus["note"] = "back-to-back device-resident steps, no L2 flush, against 4 x MEASURED_PEAKS.bf16_tflops_sustained"
        out["sustained"] = sus
    if a.latency_calls > 0 and rank == 0:
        out["latency_batch1"] = measure_latency(L, rt, fam, host_ptr, a.latency_calls)
    if world > 1:
        barrier()
    if a.faults > 0 and fam == "cnv":
        for ptr in (host_ptr, cls_ptr, words_ptr):
            native.free_pinned(ptr)
        host_ptr = None
        rt.release()
        out["faults"] = measure_faults(native, L, a.net, local_rank, rank, world, a.fault_images, a.faults, barrier, max_over_ranks)
    if a.nets and rank == 0 and world == 1:
        others = [n for n in PARAMS if n != a.net] if a.nets == "all" else [n for n in a.nets.split(",") if n in PARAMS and n != a.net]
        rt.release()
        out["networks"] = {n: measure_network(native, L, n, local_rank, 5, B) for n in others}
    if world > 1 and os.environ.get("BENCH_ONE_PROCESS", "1") != "0":
        # the engine's own multi-device path: ONE process feeding every GPU (one feeder thread per device), measured by rank 0 while the
        # other ranks wait at the barrier
        rt.release()
        barrier()
        # the idle ranks wait on the HOST (rendezvous store): an NCCL barrier would park a spinning kernel on their GPUs, which rank 0's
        # engine is about to use
        from datetime import timedelta
        store = dist.distributed_c10d._get_default_store()
        if rank == 0:
            try:
                out["e2e_one_process"] = one_process_e2e(native, L, a.net, world, B, max(a.steps // 2, 3))
            except Exception as e:
                out["e2e_one_process"] = {"error": str(e)}
            store.set("bnnb_one_process_done", "1")
        else:
            store.wait(["bnnb_one_process_done"], timedelta(minutes=15))
        barrier()
    if rank == 0 and world > 1:
        try:
            topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
            out["e2e"]["topology"] = [ln for ln in topo.splitlines() if ln.startswith("GPU") or ln.strip().startswith("GPU0")][: world + 1]
        except Exception:
            pass
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        os.sched_setaffinity(0, all_cpus)  # the CPU baselines use every host core
        try:
            ips, _, cores, sample = run_reference_cpu(a.net, 4, 1, 1)
            out["cpu_baseline"] = {"value": ips, "unit": "images/s", "cores": cores, "kind": "reference", "sample": sample}
        except Exception as e:  # the checker is missing: say so, never substitute
            out["cpu_baseline"] = {"value": None, "unit": "images/s", "cores": 0, "kind": "reference", "sample": "unavailable: %s" % e}
        try:
            out["cpu_baseline_fast"] = fast_cpu_port(a.net)
        except Exception as e:
            out["cpu_baseline_fast"] = {"value": None, "kind": "port", "sample": "unavailable: %s" % e}
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
