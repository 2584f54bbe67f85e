"""Generates tests/golden/exporter_ref.npz: outputs of the REFERENCE exporter's own batch-norm folding functions - quantize,
makeFCBNComplex, makeFCBNComplex_QNN, makeConvBNComplex (bnn/src/training/finnthesizer.py:149-330) - on seeded random batch-norm
parameters.  Run in the build container only:  python tests/golden/make_golden_exporter.py
The reference file is Python 2 (print statements, range() used as a list, tab/space mixed indentation) and imports nothing but
numpy for these functions, so the function block is taken from the file where it lies, put through three mechanical rewrites
(tabs -> 8 columns, `print "..."` lines dropped, `x = range(n)` -> `x = list(range(n))`) and executed as is: the arithmetic - the
batch-norm algebra, int() / floor / ceil, the sign flip, the binarisation - is the reference's, character for character."""
import os
import re

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("BNN_REFERENCE", "/root/reference") + "/bnn/src/training/finnthesizer.py"


def reference_functions():
    lines = open(REF).read().expandtabs(8).split("\n")[148:330]          # padTo .. end of makeConvBNComplex
    src = []
    for ln in lines:
        if re.match(r"\s*print\s+\"", ln):
            continue
        src.append(re.sub(r"=\s*range\(([^)]*)\)\s*$", r"= list(range(\1))", ln))
    ns = {"np": np}
    exec(compile("\n".join(src), REF, "exec"), ns)
    return ns


def bn_params(rng, outs, dtype):
    """Archive-like batch-norm parameters; every other neuron has a negative slope (the flip branch)."""
    gamma = rng.uniform(0.3, 2.0, outs) * np.where(np.arange(outs) % 2, -1.0, 1.0)
    return (rng.normal(0, 2.0, outs).astype(dtype), rng.normal(0, 1.0, outs).astype(dtype), gamma.astype(dtype),
            rng.normal(0, 6.0, outs).astype(dtype), rng.uniform(0.05, 1.5, outs).astype(dtype))   # bias, beta, gamma, mean, invstd


def cases():
    """(kind, name, kwargs): the parameter combinations the five networks' drivers reach (readFCBNComplex :360-372, readConvBNComplex
    :437-442; *-gen-weights-*.py), each on float32 (what an .npz archive holds) and float64 parameters."""
    out = []
    for dt in ("float32", "float64"):
        out += [("fc", "fc_w1a1_" + dt, dict(w_int=1, w_fract=0, popcount=True, dt=dt)),           # lfcW1A1, CNV W1A1 FC layers
                ("fc", "fc_signed_" + dt, dict(w_int=1, w_fract=0, popcount=False, dt=dt)),        # 1-bit output, wider input
                ("fc", "fc_w2_signed_" + dt, dict(w_int=2, w_fract=0, popcount=False, dt=dt)),
                ("fcq", "fcq_w1a2_" + dt, dict(w_int=1, w_fract=0, a_int=2, a_fract=0, tb=16, tib=None, dt=dt)),   # lfcW1A2, cnvW1A2 FC
                ("fcq", "fcq_w2a2_" + dt, dict(w_int=2, w_fract=0, a_int=2, a_fract=0, tb=16, tib=None, dt=dt)),   # cnvW2A2 FC
                ("conv", "conv_w1a1_pop_" + dt, dict(w_int=1, w_fract=0, a_int=1, a_fract=0, popcount=True, tb=16, tib=None, inter=True, dt=dt)),
                ("conv", "conv_w1a1_l0_" + dt, dict(w_int=1, w_fract=0, a_int=1, a_fract=0, popcount=False, tb=24, tib=16, inter=True, dt=dt)),  # 1/256 units
                ("conv", "conv_w1a2_" + dt, dict(w_int=1, w_fract=0, a_int=2, a_fract=0, popcount=False, tb=16, tib=None, inter=True, dt=dt)),
                ("conv", "conv_w2a2_l0_" + dt, dict(w_int=2, w_fract=0, a_int=2, a_fract=0, popcount=False, tb=24, tib=16, inter=True, dt=dt)),
                ("conv", "conv_w2a2_planar_" + dt, dict(w_int=2, w_fract=0, a_int=2, a_fract=0, popcount=False, tb=16, tib=None, inter=False, dt=dt))]
    return out


def inputs(kind, name, kw):
    # crc-free, order-independent seed: position in the case list
    rng = np.random.default_rng(1000 + [c[1] for c in cases()].index(name))
    dt = np.dtype(kw["dt"])
    if kind == "conv":
        ofm, ifm, k = 12, 5, 3
        w = rng.normal(0, 0.8, (ofm, ifm, k, k)).astype(dt)
        w.reshape(-1)[::7] = 0.0                                   # exact zeros: binarize(-0.0) on flipped neurons
        w.reshape(-1)[3::11] = 0.5                                 # round-half cases of the 2-bit quantiser
        return (w,) + bn_params(rng, ofm, dt)
    ins, outs = 40, 14
    w = rng.normal(0, 0.8, (ins, outs)).astype(dt)
    w.reshape(-1)[::7] = 0.0
    w.reshape(-1)[3::11] = -0.5
    return (w,) + bn_params(rng, outs, dt)


if __name__ == "__main__":
    ref = reference_functions()
    out = {}
    for kind, name, kw in cases():
        w, bias, beta, gamma, mean, invstd = inputs(kind, name, kw)
        if kind == "fc":
            W, T = ref["makeFCBNComplex"](w, bias, beta, gamma, mean, invstd, kw["w_int"], kw["w_fract"], use_rowmajor=True, usePopCount=kw["popcount"])
        elif kind == "fcq":
            W, T = ref["makeFCBNComplex_QNN"](w, bias, beta, gamma, mean, invstd, kw["w_fract"], kw["a_fract"], kw["w_int"], kw["a_int"], True, kw["tb"], kw["tib"])
        else:
            W, T = ref["makeConvBNComplex"](w, bias, beta, gamma, mean, invstd, kw["inter"], kw["w_fract"], kw["a_fract"], 0, kw["w_int"], kw["a_int"], 8,
                                            usePopCount=kw["popcount"], numThresBits=kw["tb"], numThresIntBits=kw["tib"])
        out[name + "/W"] = np.asarray(W, np.float64)
        out[name + "/T"] = np.asarray([np.asarray(t, np.float64) for t in T], np.float64)
    q = ref["quantize"]
    xs = np.array([-2.5, -1.5, -1.0, -0.75, -0.5, -0.25, -0.0, 0.0, 0.25, 0.5, 0.75, 1.0, 1.49, 1.5, 2.5])
    out["quantize/x"] = xs
    for (i, f) in ((1, 0), (2, 0), (1, 1), (2, 2)):
        out["quantize/%d_%d" % (i, f)] = np.array([float(q(x, i, f)) for x in xs])
    np.savez_compressed(os.path.join(HERE, "exporter_ref.npz"), **out)
    print("wrote exporter_ref.npz with %d arrays" % len(out))
