"""ctypes door to oracle/_ref/python_sw-<net>.so: the UNMODIFIED reference sw runtime (built by
oracle/build_ref.sh against stand-in HLS headers).  TEST INFRASTRUCTURE ONLY.  The reference keeps all state in
process globals, so use one network per process (run_ref_job is the subprocess entry)."""
import ctypes as C
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def ref_so(net):
    return os.path.join(ROOT, "oracle", "_ref", "python_sw-%s.so" % net)


class Ref:
    def __init__(self, net, pdir):
        self.net = net
        self.is_cnv = net.startswith("cnv")
        L = self.L = C.CDLL(ref_so(net))
        L.load_parameters.argtypes = [C.c_char_p]
        L.inference.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p]
        L.inference_multiple.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.inference_multiple.restype = C.POINTER(C.c_int)
        L.free_results.argtypes = [C.POINTER(C.c_int)]
        L.ref_inject_fault.argtypes = [C.c_int, C.c_int, C.c_uint, C.c_uint]
        L.ref_mem_read.argtypes = [C.c_uint] * 4
        L.ref_mem_read.restype = C.c_uint64
        L.ref_mem_set.argtypes = [C.c_uint] * 4 + [C.c_uint64]
        L.ref_blackboxjam.argtypes = [C.c_void_p, C.c_void_p, C.c_uint]
        L.load_parameters(pdir.encode())

    def _file(self, data):
        f = tempfile.NamedTemporaryFile(delete=False, suffix=".bin")
        f.write(data)
        f.close()
        return f.name

    def multi(self, filebytes, ncls=10, detail=0):
        path = self._file(filebytes)
        n = C.c_int(0)
        us = C.c_float(0)
        r = self.L.inference_multiple(path.encode(), ncls, C.byref(n), C.byref(us), detail)
        cnt = n.value * (ncls if (detail and self.is_cnv) else 1)
        out = np.array(r[:cnt], np.int32)
        self.L.free_results(r)
        os.unlink(path)
        return out.reshape(n.value, -1) if (detail and self.is_cnv) else out

    def single(self, filebytes, ncls=10):
        path = self._file(filebytes)
        res = (C.c_int * 64)()
        us = C.c_float(0)
        cls = self.L.inference(path.encode(), res, ncls, C.byref(us))
        os.unlink(path)
        return cls, np.array(res[:64], np.int32)

    def inject(self, target, layer, bit, word_size=1):
        """target: 0 weights, 1 thresholds (the C ABI's convention, main_python.cpp:105-108)."""
        self.L.ref_inject_fault(target, word_size, layer, bit)

    def mem_read(self, tl, mem, ind, th):
        return self.L.ref_mem_read(tl, mem, ind, th)

    def mem_set(self, tl, mem, ind, th, val):
        self.L.ref_mem_set(tl, mem, ind, th, val & 0xFFFFFFFFFFFFFFFF)

    def raw(self, packed_in, count, out_words):
        """BlackBoxJam compute mode on already packed inputs (uint64)."""
        packed_in = np.ascontiguousarray(packed_in, np.uint64)
        out = np.zeros(count * out_words, np.uint64)
        self.L.ref_blackboxjam(packed_in.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), count)
        return out.reshape(count, out_words)
