#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_norm_gpu.py tests/test_nvfp4_gpu.py tests/test_gemm_gpu.py tests/test_gemm_fullsize_gpu.py tests/test_guard_gpu.py tests/test_linear_gpu.py -x -q -m gpu 2>&1 | tail -4 | cut -c1-300
bash scripts/gpu_kernel_bench.sh --only gemm_fp8 | tail -5
for bn in 256 224 192; do echo "== forced BN $bn"; NVTE_B200_GEMM_BN=$bn bash scripts/gpu_kernel_bench.sh --only gemm_fp8 | tail -4; done
