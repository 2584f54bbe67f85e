#!/bin/bash
# on the GPU box: the whole GPU suite, then one bench line with the side measurements given as arguments
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_tests.log
python bench.py --steps 10 --warmup 3 "$@" > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err
tail -4 gpurun_out/r2_tests.log; tail -3 gpurun_out/r2_bench.err
