#!/bin/bash
# on the GPU box: launch list of the plain bench command, then one full capture of the path's kernels (B200_PROFILING.md recipe)
set -x
python bench.py --steps 2 --warmup 1 --lean > gpurun_out/r2_ncu_plain.json 2> gpurun_out/r2_ncu_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 1 --lean > gpurun_out/r2_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_cnv01|k_conv_f4|k_gemm_f4|k_fc_chain|k_cnv_classes' -s 18 -c 6 -f -o gpurun_out/r2_prof \
    python bench.py --steps 2 --warmup 1 --lean > gpurun_out/r2_ncu_f.log 2>&1
ls -la gpurun_out/r2_prof.ncu-rep
