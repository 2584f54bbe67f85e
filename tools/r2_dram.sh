#!/bin/bash
# on the GPU box: DRAM bytes of the path kernels over 4 device-resident steps of 65 536 images at several chunk sizes, L2 left alone
for c in 65536 16384 8192 4096; do
  BNN_B200_CHUNK=$c ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --cache-control none --clock-control none \
    -k regex:'k_cnv01|k_conv_f4|k_gemm_f4|k_fc_chain|k_cnv_classes' --csv --log-file gpurun_out/r2_dram_$c.csv \
    python tools/dev_step.py cnvW1A1 65536 4 > gpurun_out/r2_dram_$c.log 2>&1
done
