#!/bin/bash
# on the GPU box: tools/run_variants.sh v1 v2 ... -> bench line summary (+ PROF01 lines if the variant prints them)
for v in "$@"; do
  BNNB_BENCH_NOCHECK=$NOCHECK BNN_B200_LIB_DIR=$PWD/variants/$v python bench.py --steps 5 --warmup 3 --lean > gpurun_out/v_$v.out 2> gpurun_out/v_$v.err
  echo "== $v"
  grep PROF01 gpurun_out/v_$v.out | grep -E "warp (0|8|16|18) " | tail -4 | cut -c1-260
  python - <<PY
import json
for l in open("gpurun_out/v_$v.out"):
    if l.startswith("{"):
        d = json.loads(l); print("value %.2f M  e2e %.2f M  layer_ms %s" % (d["value"]/1e6, d["e2e"]["value"]/1e6, [round(x,3) for x in d["roofline"]["layer_ms_per_step"]]))
PY
done
