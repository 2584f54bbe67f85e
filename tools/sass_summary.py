#!/usr/bin/env python
"""Blackwell evidence for the hot kernels, from the in-tree build (the .so is git-ignored, so the listing is committed instead):
   tools/sass_summary.py [round-tag]  ->  profiles/<tag>_sass_summary.txt
For every hot kernel of lib/libbnn_b200.so (production instantiations, i.e. not the TAP ones): ptxas registers / shared memory /
spills (build/*.ptxas.log), the count of each tensor-core, TMEM, TMA and mbarrier SASS mnemonic (cuobjdump -sass), and the first
lines on which UTCOMMA / UTCIMMA / LDTM / STTM / UBLKCP / UTCBAR appear."""
import os
import re
import subprocess
import sys
from collections import Counter, OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "bnn-pynq_b200")
HOT = ("k_cnv01_tc", "k_conv_f4", "k_gemm_f4", "k_fc_chain", "k_conv_tc", "k_gemm_tc", "k_layer_popc", "k_conv0", "k_patch_apply", "k_prep_cifar", "k_prep_mnist")
MNEMONICS = ("UTCOMMA", "UTCOMMA.2CTA", "UTCBAR.2CTA", "UTCIMMA", "UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "SYNCS", "POPC", "LOP3", "SHF", "PRMT", "IMMA", "IDP")
SHOW = ("UTCOMMA", "UTCIMMA", "LDTM", "STTM", "UBLKCP", "UTCBAR")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def ptxas_info():
    info = {}
    for f in sorted(os.listdir(os.path.join(PKG, "build"))):
        if not f.endswith(".ptxas.log"):
            continue
        cur = None
        for ln in open(os.path.join(PKG, "build", f)):
            m = re.search(r"Compiling entry function '([^']+)'", ln)
            if m:
                cur = m.group(1)
                info[cur] = []
            elif cur and ("Used" in ln or "spill" in ln):
                info[cur].append(ln.strip().replace("ptxas info    : ", ""))
    return info


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(PKG, "lib", "libbnn_b200.so")], capture_output=True, text=True).stdout
    funcs = OrderedDict()
    cur = None
    for ln in sass.split("\n"):
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
            funcs[cur].append(ln.strip())
    dm = demangle(list(funcs))
    pinfo = ptxas_info()
    lines = ["# SASS / ptxas summary of bnn-pynq_b200/lib/libbnn_b200.so (sm_100a), %s; written by tools/sass_summary.py" % tag, ""]
    total = Counter()
    for mangled, body in funcs.items():
        name = dm.get(mangled, mangled)
        if not any(h + "<" in name or h + "(" in name for h in HOT):
            continue
        # TAP flag: the last template argument of k_cnv01_tc / k_gemm_f4, the last but one of k_conv_f4 (whose last is PAIR)
        m = re.search(r"((?:\(bool\)[01]|true|false)(?:, (?:\(bool\)[01]|true|false))?)>\(", name)
        flags = [f in ("(bool)1", "true") for f in m.group(1).split(", ")] if m else []
        tap = bool(flags) and (flags[0] if ("k_conv_f4" in name and len(flags) == 2) else flags[-1]) and ("k_cnv01_tc" in name or "k_conv_f4" in name or "k_gemm_f4" in name)
        cnt = Counter()
        first = {}
        for i, ins in enumerate(body):
            for mn in MNEMONICS:
                if (mn in ins) if "." in mn else re.search(r"\b%s\b|\b%s\." % (mn, mn), ins):
                    cnt[mn] += 1
                    first.setdefault(mn, ins)
        if not tap:
            total.update(cnt)
        short = re.sub(r"bnnb::\(anonymous namespace\)::|bnnb::", "", name)
        short = re.sub(r"\(bnnb::DevLayer.*", "", short)
        lines.append("## %s%s" % (short, "   [TAP instantiation: parity tests only]" if tap else ""))
        lines.append("instructions: %d" % len(body))
        for p in pinfo.get(mangled, []):
            lines.append("ptxas: " + p)
        lines.append("mnemonics: " + ", ".join("%s %d" % (k, cnt[k]) for k in MNEMONICS if cnt[k]))
        if not tap:
            for mn in SHOW:
                if mn in first:
                    lines.append("  first %-8s %s" % (mn, first[mn][:150]))
        lines.append("")
    lines.insert(2, "totals over the production instantiations: " + ", ".join("%s %d" % (k, total[k]) for k in MNEMONICS if total[k]))
    lines.insert(3, "")
    out = os.path.join(ROOT, "profiles", "%s_sass_summary.txt" % tag)
    open(out, "w").write("\n".join(lines))
    print("wrote", out, "(%d kernels)" % sum(1 for l in lines if l.startswith("## ")))


if __name__ == "__main__":
    main()
