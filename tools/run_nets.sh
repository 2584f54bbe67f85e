#!/bin/bash
# on the GPU box: tools/run_nets.sh <variant|-> <net>... -> one summary line per network ("-" = the in-tree build)
v=$1; shift
for n in "$@"; do
  if [ "$v" = "-" ]; then unset BNN_B200_LIB_DIR; else export BNN_B200_LIB_DIR=$PWD/variants/$v; fi
  python bench.py --net $n --steps 10 --warmup 3 --lean > gpurun_out/n_${v}_$n.out 2> gpurun_out/n_${v}_$n.err
  python - <<PY
import json
for l in open("gpurun_out/n_${v}_$n.out"):
    if l.startswith("{"):
        d = json.loads(l); print("$v $n value %.2f M  e2e %.2f M  layer_ms %s" % (d["value"]/1e6, d["e2e"]["value"]/1e6, [round(x,3) for x in d["roofline"]["layer_ms_per_step"]]))
PY
done
