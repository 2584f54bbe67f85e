import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "bnn-pynq_b200"))
from bnn import native
names = {0: "popc", 2: "lop3", 3: "shf.l.wrap", 4: "prmt", 5: "imad.hi", 6: "imad", 7: "shr", 8: "fmnmx"}
for w, n in names.items():
    r, clk = native.microbench(w)
    print("%-12s %.3e /s  clock %.0f MHz  -> %.1f thread-instr/clk/SM" % (n, r, clk, r / (148 * 1965e6)))
