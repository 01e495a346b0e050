#!/bin/bash
mkdir -p gpurun_out
python scripts/ncu_target.py mxfp8 4 > gpurun_out/plain_mx.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:quantize_tma_kernel -s 2 -c 1 -o gpurun_out/r02_prof_mxfp8_dy -f python scripts/ncu_target.py mxfp8 4 > gpurun_out/ncu_mx.log 2>&1
tail -3 gpurun_out/ncu_mx.log
