# A/B of the fc_chain kernel against per-layer launches (device-resident steps, per-layer CUDA-event times)
for net in ${NETS:-cnvW1A1 cnvW1A2 lfcW1A1 lfcW1A2}; do
  for c in 0 1; do
    BNN_B200_CHAIN=$c DEV_STEP_PROFILE=chain$c timeout 200 python tools/dev_step.py $net ${B:-65536} 8 2>&1 | grep -E "layer_ms|rror"
  done
done
