#!/bin/bash
# on the GPU box: one full ncu capture of a step on the kind::i8 core and one on the LOP3+POPC core, beside the fp4 capture of
# tools/r2_ncu.sh (the evidence behind "core chosen by measurement"); plain runs first (numbers under ncu are never bench values)
set -x
for c in 1 0; do
  b=65536; [ $c = 0 ] && b=8192
  python bench.py --core $c --batch $b --steps 2 --warmup 1 --lean > gpurun_out/r2_core${c}_plain.json 2> gpurun_out/r2_core${c}_plain.err || exit 1
  ncu --set full --clock-control none -k regex:'k_cnv01|k_conv_tc|k_gemm_tc|k_conv_f4|k_gemm_f4|k_layer_popc|k_conv0|k_cnv_classes' -s 27 -c 9 -f \
      -o gpurun_out/r2_core$c python bench.py --core $c --batch $b --steps 2 --warmup 1 --lean > gpurun_out/r2_core${c}_ncu.log 2>&1
done
ls -la gpurun_out/r2_core*.ncu-rep
