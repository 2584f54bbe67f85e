"""ctypes binding of the handle-based C ABI (include/bnn_b200.h, section 2) exported by lib/libbnn_b200.so.
In-memory entry points for callers that already hold the images (no temp-file round trip), deterministic fault lists,
accelerator-memory access and device-resident inference.  There is no CPU fallback: every call needs the built library."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.realpath(__file__))
LIB_DIR = os.environ.get("BNN_B200_LIB_DIR", os.path.join(os.path.dirname(_HERE), "lib"))
_lib = None

_SIGS = {
    "bnnb_open": (C.c_void_p, [C.c_char_p]),
    "bnnb_close": (None, [C.c_void_p]),
    "bnnb_last_error": (C.c_char_p, [C.c_void_p]),
    "bnnb_set_devices": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "bnnb_num_devices": (C.c_int, [C.c_void_p]),
    "bnnb_load_parameters": (C.c_int, [C.c_void_p, C.c_char_p]),
    "bnnb_release": (C.c_int, [C.c_void_p]),
    "bnnb_device_bytes": (C.c_size_t, [C.c_void_p, C.c_int]),
    "bnnb_classify_cifar": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "bnnb_classify_mnist": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "bnnb_classify_packed": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "bnnb_classify_with_fault_list": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "bnnb_schedule_faults": (C.c_int, [C.c_void_p, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_void_p, C.c_uint, C.c_longlong, C.c_void_p, C.c_int]),
    "bnnb_inject_fault": (C.c_int, [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_uint]),
    "bnnb_mem_read": (C.c_uint64, [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_uint]),
    "bnnb_mem_set": (C.c_int, [C.c_void_p, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_uint64]),
    "bnnb_weight_layer_bits": (C.c_uint, [C.c_void_p, C.c_int]),
    "bnnb_thresh_layer_bits": (C.c_uint, [C.c_void_p, C.c_int]),
    "bnnb_tap_acc_size": (C.c_size_t, [C.c_void_p]),
    "bnnb_tap_act_size": (C.c_size_t, [C.c_void_p]),
    "bnnb_tap": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "bnnb_infer_device": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "bnnb_infer_device_static": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]),
    "bnnb_prep_device": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "bnnb_sync": (C.c_int, [C.c_void_p, C.c_int]),
    "bnnb_last_launches": (C.c_int, [C.c_void_p]),
    "bnnb_in_bytes": (C.c_int, [C.c_void_p]),
    "bnnb_out_bytes": (C.c_int, [C.c_void_p]),
    "bnnb_set_core": (C.c_int, [C.c_void_p, C.c_int]),
    "bnnb_profile_enable": (C.c_int, [C.c_void_p, C.c_int]),
    "bnnb_profile_collect": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "bnnb_microbench": (C.c_double, [C.c_int, C.c_void_p]),
    "bnnb_tc_probe_fp4": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "bnnb_tc_probe": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "bnnb_alloc_host": (C.c_void_p, [C.c_size_t]),
    "bnnb_free_host": (None, [C.c_void_p]),
}


def lib_path():
    return os.path.join(LIB_DIR, "libbnn_b200.so")


def lib():
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise RuntimeError("%s is missing: build it with `make -C bnn-pynq_b200`; this runtime has no CPU fallback" % p)
        L = C.CDLL(p)
        for name, (res, args) in _SIGS.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Runtime:
    """One network on the visible GPUs."""

    def __init__(self, network, params_dir=None, devices=None):
        self.L = lib()
        self.network = network
        self.is_cnv = network.startswith("cnv")
        self.h = self.L.bnnb_open(network.encode())
        if not self.h:
            raise ValueError("unknown network %r" % network)
        self._devices = None if devices is None else [int(d) for d in devices]
        if devices is not None:
            ids = np.ascontiguousarray(devices, np.int32)
            self._ck(self.L.bnnb_set_devices(self.h, _p(ids), len(ids)))
        if params_dir is not None:
            self.load_parameters(params_dir)

    def close(self):
        if self.h:
            self.L.bnnb_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError("bnn_b200[%s]: %s" % (self.network, self.L.bnnb_last_error(self.h).decode()))

    def load_parameters(self, params_dir):
        self._ck(self.L.bnnb_load_parameters(self.h, params_dir.encode()))

    def release(self):
        """Free every device allocation; parameters and faults stay in the host mirror (the next call rebuilds the device state)."""
        self._ck(self.L.bnnb_release(self.h))

    def device_bytes(self, slot=0):
        return self.L.bnnb_device_bytes(self.h, slot)

    @property
    def usec_per_image(self):
        return self._usec

    def classify_cifar(self, records, number_class=10, want_scores=False):
        """records: uint8 [n, 3073].  Returns classes (and int16 scores [n, 64])."""
        rec = np.ascontiguousarray(records, np.uint8).reshape(-1, 3073)
        n = rec.shape[0]
        cls = np.zeros(n, np.int32)
        sc = np.zeros((n, 64), np.int16) if want_scores else None
        us = C.c_float(0)
        self._ck(self.L.bnnb_classify_cifar(self.h, _p(rec), n, number_class, _p(cls), _p(sc), C.byref(us)))
        self._usec = us.value
        return (cls, sc) if want_scores else cls

    def classify_mnist(self, pixels, number_class=10, want_words=False):
        px = np.ascontiguousarray(pixels, np.uint8).reshape(-1, 784)
        n = px.shape[0]
        cls = np.zeros(n, np.int32)
        words = np.zeros(n, np.uint64) if want_words else None
        us = C.c_float(0)
        self._ck(self.L.bnnb_classify_mnist(self.h, _p(px), n, number_class, _p(cls), _p(words), C.byref(us)))
        self._usec = us.value
        return (cls, words) if want_words else cls

    def classify_packed(self, packed, number_class=10, want_raw=False):
        """Inputs already in BlackBoxJam's format: uint8 [n, 104] (LFC, see pack_mnist) or int8 [n, 3072] HWC (CNV).  Returns classes
        (and with want_raw the int16 scores [n, 64] / the uint64 output words)."""
        a = np.ascontiguousarray(packed).view(np.uint8).reshape(-1, self.in_bytes)
        n = a.shape[0]
        cls = np.zeros(n, np.int32)
        sc = np.zeros((n, 64), np.int16) if (want_raw and self.is_cnv) else None
        words = np.zeros(n, np.uint64) if (want_raw and not self.is_cnv) else None
        us = C.c_float(0)
        self._ck(self.L.bnnb_classify_packed(self.h, _p(a), n, number_class, _p(cls), _p(sc), _p(words), C.byref(us)))
        self._usec = us.value
        return (cls, sc if self.is_cnv else words) if want_raw else cls

    def classify(self, data, number_class=10):
        return self.classify_cifar(data, number_class) if self.is_cnv else self.classify_mnist(data, number_class)

    def classify_with_fault_list(self, data, faults, number_class=10):
        w = 3073 if self.is_cnv else 784
        d = np.ascontiguousarray(data, np.uint8).reshape(-1, w)
        fl = np.ascontiguousarray(faults, np.uint32).reshape(-1, 5)
        cls = np.zeros(d.shape[0], np.int32)
        us = C.c_float(0)
        self._ck(self.L.bnnb_classify_with_fault_list(self.h, _p(d), d.shape[0], number_class, _p(fl), fl.shape[0], _p(cls), C.byref(us)))
        self._usec = us.value
        return cls

    def schedule_faults(self, num_images, flip_count, word_size=1, target=-1, target_layers=None, seed=-1):
        out = np.zeros((max(flip_count, 1), 5), np.uint32)
        tl = None if not target_layers else np.ascontiguousarray(target_layers, np.int32)
        m = self.L.bnnb_schedule_faults(self.h, num_images, flip_count, word_size, target, _p(tl), 0 if tl is None else len(tl), seed, _p(out), flip_count)
        if m < 0:
            self._ck(m)
        return out[:m]

    def inject_fault(self, target, layer, bit, word_size=1):
        self._ck(self.L.bnnb_inject_fault(self.h, target, layer, bit, word_size))

    def mem_read(self, tl, mem, ind, th):
        return self.L.bnnb_mem_read(self.h, tl, mem, ind, th)

    def mem_set(self, tl, mem, ind, th, val):
        self._ck(self.L.bnnb_mem_set(self.h, tl, mem, ind, th, val & 0xFFFFFFFFFFFFFFFF))

    def weight_layer_bits(self, l):
        return self.L.bnnb_weight_layer_bits(self.h, l)

    def thresh_layer_bits(self, l):
        return self.L.bnnb_thresh_layer_bits(self.h, l)

    def tap(self, data):
        w = 3073 if self.is_cnv else 784
        d = np.ascontiguousarray(data, np.uint8).reshape(-1, w)
        acc = np.zeros((d.shape[0], self.L.bnnb_tap_acc_size(self.h)), np.int32)
        act = np.zeros((d.shape[0], self.L.bnnb_tap_act_size(self.h)), np.int8)
        self._ck(self.L.bnnb_tap(self.h, _p(d), d.shape[0], _p(acc), _p(act)))
        return acc, act

    # device-resident path (pointers are integers, e.g. torch.Tensor.data_ptr())
    def infer_device(self, d_in, n, d_out, d_classes=0, number_class=10, stream=0, slot=0):
        self._ck(self.L.bnnb_infer_device(self.h, slot, d_in, n, d_out, d_classes or None, number_class, stream or None))

    def infer_device_static(self, d_in, n, d_out, d_classes=0, number_class=10, stream=0, slot=0):
        """infer_device without any allocation: activation scratch in the library's static 32 MiB device array (include/bnn_b200.h)."""
        self._ck(self.L.bnnb_infer_device_static(self.h, slot, d_in, n, d_out, d_classes or None, number_class, stream or None))

    def prep_device(self, d_raw, n, d_in, stream=0, slot=0):
        self._ck(self.L.bnnb_prep_device(self.h, slot, d_raw, n, d_in, stream or None))

    def sync(self, slot=0):
        self._ck(self.L.bnnb_sync(self.h, slot))

    def classify_tensor(self, images, number_class=10, want_scores=False):
        """Device-resident entry point: `images` is a CUDA uint8 tensor (or any DLPack exporter on a CUDA device) of CIFAR records
        [n, 3073] (CNV) or pixels [n, 784] (LFC).  Runs the quantise/pack kernel and the network on the tensor's device, ordered
        after torch's current stream, and returns an int32 class tensor on that device (CNV with want_scores: also int16 scores [n, 64]; LFC with
        want_scores: also the raw 64-bit output words as int64).  Nothing crosses PCIe."""
        import torch
        t = images if isinstance(images, torch.Tensor) else torch.from_dlpack(images)
        if not t.is_cuda or t.dtype != torch.uint8:
            raise ValueError("classify_tensor needs a CUDA uint8 tensor (host arrays go through classify_cifar / classify_mnist)")
        w = 3073 if self.is_cnv else 784
        t = t.contiguous().reshape(-1, w)
        n = t.shape[0]
        dev = t.device
        slot = dev.index if self._devices is None else self._devices.index(dev.index)
        with torch.cuda.device(dev):
            # torch's default stream is the NULL handle, which the C ABI reads as "the engine's own stream": run on a side stream that
            # is ordered after the caller's current stream, and make the current stream wait for it before returning
            cur = torch.cuda.current_stream()
            if not hasattr(self, "_side"):
                self._side = {}
            side = self._side.setdefault(dev.index, torch.cuda.Stream(dev))
            stream = side.cuda_stream
            d_in = torch.empty(max(n, 1) * self.in_bytes, dtype=torch.uint8, device=dev)
            d_out = torch.empty(max(n, 1) * self.out_bytes, dtype=torch.uint8, device=dev)
            cls = torch.zeros(n, dtype=torch.int32, device=dev)
            if n == 0:
                return (cls, d_out[:0].view(torch.int16).reshape(0, 64) if self.is_cnv else d_out[:0].view(torch.int64)) if want_scores else cls
            side.wait_stream(cur)                        # after the allocations / zero-fill above, which run on `cur`
            self.prep_device(t.data_ptr(), n, d_in.data_ptr(), stream, slot)
            self.infer_device(d_in.data_ptr(), n, d_out.data_ptr(), cls.data_ptr() if self.is_cnv else 0, number_class, stream, slot)
            cur.wait_stream(side)
            if self.is_cnv:
                return (cls, d_out.view(torch.int16).reshape(n, 64)) if want_scores else cls
            # LFC multi-image rule (foldedmv-offload.cpp:202-220): (unsigned) log2((double)(word & mask)), 0 when no bit is set
            words = d_out.view(torch.int64)
            nc = max(0, min(64, number_class))
            mask = -1 if nc == 64 else (1 << nc) - 1
            m = words & mask
            as_double = ((m >> 32) & 0xFFFFFFFF).to(torch.float64) * 4294967296.0 + (m & 0xFFFFFFFF).to(torch.float64)
            # floor(log2) as the binary exponent of the rounded double: equal to glibc's log2 for every word whose top 48 bits below
            # the leading one are not all ones (log2 of such a word rounds up to the next integer on the host); padded LFC rows carry
            # threshold 32767 and never fire, so those words need more than 48 real classes
            cls = torch.where(m != 0, torch.frexp(as_double)[1] - 1, torch.zeros_like(m, dtype=torch.int32)).to(torch.int32)
            return (cls, words) if want_scores else cls

    def set_core(self, core):
        """0: LOP3+POPC everywhere; 1: tcgen05 tensor-core kernels where instantiated."""
        self.L.bnnb_set_core(self.h, core)

    def profile_enable(self, on=True):
        self.L.bnnb_profile_enable(self.h, 1 if on else 0)

    def profile_collect(self):
        """-> (ms[10], launches[10]): layers 0..8, class decode."""
        ms = np.zeros(10, np.float32)
        cnt = np.zeros(10, np.int32)
        self.L.bnnb_profile_collect(self.h, _p(ms), _p(cnt))
        return ms, cnt

    @property
    def last_launches(self):
        return self.L.bnnb_last_launches(self.h)

    @property
    def in_bytes(self):
        return self.L.bnnb_in_bytes(self.h)

    @property
    def out_bytes(self):
        return self.L.bnnb_out_bytes(self.h)


def pack_mnist(pixels):
    """binarizeAndPack on the host (foldedmv-offload.cpp:82-98): uint8 [n, 784] -> uint8 [n, 104]; bit k = (pixel k >= 128), LSB first,
    784 -> 832 bits zero padded."""
    px = np.ascontiguousarray(pixels, np.uint8).reshape(-1, 784)
    bits = np.zeros((px.shape[0], 832), np.uint8)
    bits[:, :784] = px >= 128
    return np.ascontiguousarray(np.packbits(bits, axis=1, bitorder="little"))


def microbench(which):
    """which: 0 POPC/s, 1 int8 mma.sync MAC/s, 2 LOP3/s.  Returns (rate, implied SM clock in MHz)."""
    clk = C.c_double(0)
    r = lib().bnnb_microbench(which, C.byref(clk))
    return r, clk.value


def tc_probe(A, B, shift=0, reps=1):
    """tcgen05 kind::i8 probe: D[m][n] = sum_k A[m+shift][k] * B[n][k] for m < 128.  Returns (D int32 [128, N], cycles)."""
    A = np.ascontiguousarray(A, np.int8)
    B = np.ascontiguousarray(B, np.int8)
    D = np.zeros((128, B.shape[0]), np.int32)
    cyc = C.c_longlong(0)
    rc = lib().bnnb_tc_probe(_p(A), _p(B), _p(D), A.shape[0], B.shape[0], A.shape[1], shift, reps, C.byref(cyc))
    if rc != 0:
        raise RuntimeError("tc_probe failed: rc=%d" % rc)
    return D, cyc.value


_E2M1 = {0: 0x0, 1: 0x2, -1: 0xA, 2: 0x4, -2: 0xC}


def pack_fp4(M):
    """int matrix with values in {-2,-1,0,1,2} -> e2m1 nibbles, two per byte (even k in the low nibble)."""
    M = np.asarray(M)
    lut = np.zeros(5, np.uint8)
    for v, c in _E2M1.items():
        lut[v + 2] = c
    nib = lut[M + 2]
    return np.ascontiguousarray(nib[:, 0::2] | (nib[:, 1::2] << 4))


def tc_probe_fp4(A, B, shift=0, reps=1, bias=None, mix=0):
    """tcgen05 kind::mxf4 (block-scaled fp4, all scales 1.0) probe: D[m][n] = sum_k A[m+shift][k] * B[n][k].
    With bias (float32 [N]) the accumulator is pre-loaded with bias[n] and D is returned as float32 (raw accumulator)."""
    A4, B4 = pack_fp4(A), pack_fp4(B)
    D = np.zeros((128, B.shape[0]), np.int32)
    cyc = C.c_longlong(0)
    if mix:
        rc = lib().bnnb_tc_probe_fp4_mix(_p(A4), _p(B4), _p(D), A.shape[0], B.shape[0], A.shape[1], shift, reps, C.byref(cyc), None, mix)
    elif bias is None:
        rc = lib().bnnb_tc_probe_fp4(_p(A4), _p(B4), _p(D), A.shape[0], B.shape[0], A.shape[1], shift, reps, C.byref(cyc))
    else:
        bias = np.ascontiguousarray(bias, np.float32)
        assert bias.shape == (B.shape[0],)
        rc = lib().bnnb_tc_probe_fp4_bias(_p(A4), _p(B4), _p(D), A.shape[0], B.shape[0], A.shape[1], shift, reps, C.byref(cyc), _p(bias))
    if rc != 0:
        raise RuntimeError("tc_probe_fp4 failed: rc=%d" % rc)
    return (D.view(np.float32) if bias is not None else D), cyc.value


def pinned_array(nbytes):
    """uint8 numpy view over cudaHostAlloc'ed memory (kept alive by the returned array's base object)."""
    L = lib()
    p = L.bnnb_alloc_host(nbytes)
    if not p:
        raise MemoryError("cudaHostAlloc failed")
    buf = (C.c_uint8 * nbytes).from_address(p)
    arr = np.frombuffer(buf, np.uint8)
    return arr, p


def free_pinned(p):
    lib().bnnb_free_host(p)


def bind_host_to_device(index):
    """Pin the calling process to the CPUs (hence, by first touch, the memory node) NVML reports as local to GPU `index`.
    Call it before allocating pinned buffers: with one process per GPU the host-buffer path is PCIe- and host-memory-bound, and
    a buffer on the other socket halves its copy rate.  Returns the CPU set, or None when NVML / the affinity call is unavailable."""
    try:
        import pynvml as N
        N.nvmlInit()
        h = N.nvmlDeviceGetHandleByIndex(int(index))
        words = N.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, v in enumerate(words) for b in range(64) if (int(v) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None


def tmem_pack_probe(cols):
    """tcgen05.ld .32x32b.x16.pack::16b probe: `cols` is uint32 [32 lanes][32 columns] written to TMEM; returns the 16 packed
    registers per lane, uint32 [32][16]."""
    cols = np.ascontiguousarray(cols, np.uint32)
    assert cols.shape == (32, 32)
    out = np.zeros((32, 16), np.uint32)
    if lib().bnnb_tmem_pack_probe(_p(cols), _p(out)) != 0:
        raise RuntimeError("tmem_pack_probe failed")
    return out


def alu_rate(which):
    """Thread-instructions per second of one ALU / FMA-pipe instruction (bnnb_microbench): 0 popc, 2 lop3, 3 shf.l.wrap, 4 prmt,
    5 mad.hi, 6 mad.lo, 8 min.f32."""
    return microbench(which)[0]


def tmem_rate(warps, store=False):
    """TMEM read (tcgen05.ld) or write (tcgen05.st) throughput in bytes per clock per SM with `warps` warps issuing."""
    L = lib()
    L.bnnb_tmem_rate.restype = C.c_double
    v = L.bnnb_tmem_rate(int(warps), int(bool(store)))
    if v < 0:
        raise RuntimeError("tmem_rate failed")
    return v


def tc_rate_fp4(N, taps=False):
    """Clocks per tcgen05 kind::mxf4 MMA (M=128, K=64) at width N from a lean, fully unrolled issue loop."""
    L = lib()
    L.bnnb_tc_rate_fp4.restype = C.c_double
    v = L.bnnb_tc_rate_fp4(int(N), int(bool(taps)))
    if v < 0:
        raise RuntimeError("tc_rate_fp4 failed")
    return v


def mbar_roundtrip(mode=0):
    """Clocks per mbarrier ping-pong round trip between two warps (mode 1: one hop is a tcgen05.commit)."""
    L = lib()
    L.bnnb_mbar_roundtrip.restype = C.c_double
    return L.bnnb_mbar_roundtrip(int(mode))
