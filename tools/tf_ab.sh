for net in cnvW1A1 cnvW1A2 lfcW1A1; do
  DEV_STEP_PROFILE=v timeout 200 python tools/dev_step.py $net 65536 8 2>&1 | grep -E "layer_ms|rror"
done
timeout 600 python -m pytest tests/test_gpu_chain.py tests/test_gpu_parity.py tests/test_gpu_pairs.py -m gpu -x -q 2>&1 | tail -2
