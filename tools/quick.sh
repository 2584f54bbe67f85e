#!/bin/bash
# on the GPU box: parity tests + one bench line summary
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/quick.json 2> gpurun_out/quick.err
python - <<PY
import json
for l in open("gpurun_out/quick.json"):
    if l.startswith("{"):
        d = json.loads(l); print("value %.2f M  e2e %.2f M  layer_ms %s" % (d["value"]/1e6, d["e2e"]["value"]/1e6, [round(x,3) for x in d["roofline"]["layer_ms_per_step"]]))
PY
