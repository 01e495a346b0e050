#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_norm_gpu.py -x -q -m gpu 2>&1 | tail -2
for subs in 8 16; do
  echo "== NVTE_B200_NORM_SUBS=$subs"
  NVTE_B200_NORM_SUBS=$subs NVTE_B200_NORM_VERBOSE=1 timeout 300 python scripts/dev_norm.py 16384x8192 2>&1 | grep -v "^$" | head -14
done
