"""Per-chunk timeline of the host-buffer path (BNN_B200_TIMELINE=1): where a bnnb_classify_cifar call of 65 536 images spends its time."""
import os, sys, time
os.environ["BNN_B200_TIMELINE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bnn-pynq_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from oracle_py import params_dir
from bnn import native
net = sys.argv[1] if len(sys.argv) > 1 else "cnvW1A1"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
rt = native.Runtime(net, params_dir(net), devices=[0])
rec = 3073 if net.startswith("cnv") else 784
buf, _p = native.pinned_array(n * rec)
buf[:] = np.random.default_rng(1).integers(0, 256, n * rec, dtype=np.uint8)
img = buf.reshape(n, rec)
for it in range(4):
    if it < 3:
        fd = os.dup(2); dn = os.open(os.devnull, os.O_WRONLY); os.dup2(dn, 2)
    t = time.perf_counter()
    cls = rt.classify_cifar(img, 10) if rec == 3073 else rt.classify_mnist(img)
    dt = time.perf_counter() - t
    if it < 3:
        os.dup2(fd, 2); os.close(fd); os.close(dn)
print("call %.3f ms -> %.2f M images/s" % (dt * 1e3, n / dt / 1e6))
