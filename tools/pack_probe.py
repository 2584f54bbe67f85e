import sys, os, ctypes as C
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "bnn-pynq_b200"))
from bnn import native
L = native.lib()
rng = np.random.default_rng(0)
a = rng.integers(0, 2**32, (32, 32), dtype=np.uint64).astype(np.uint32)
out = np.zeros((32, 16), np.uint32)
rc = L.bnnb_tmem_pack_probe(a.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
print("rc", rc)
lo = a & 0xFFFF
exp = (lo[:, 0::2] | (lo[:, 1::2] << 16)).astype(np.uint32)
print("low16(col 2j) | low16(col 2j+1) << 16 :", np.array_equal(out, exp))
if not np.array_equal(out, exp):
    print([hex(x) for x in a[0, :4]], [hex(x) for x in out[0, :4]])
