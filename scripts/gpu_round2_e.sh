#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gemm_gpu.py tests/test_gemm_fullsize_gpu.py tests/test_guard_gpu.py tests/test_grouped_linear_gpu.py tests/test_linear_gpu.py -x -q -m gpu 2>&1 | tail -3 | cut -c1-300
echo "== default (PDL on, small-problem heuristic on)"; bash scripts/gpu_kernel_bench.sh --only gemm_ | tail -8
echo "== PDL off"; NVTE_B200_GEMM_PDL=0 bash scripts/gpu_kernel_bench.sh --only gemm_fp8 | tail -4
echo "== small off"; NVTE_B200_GEMM_SMALL=0 bash scripts/gpu_kernel_bench.sh --only gemm_fp8_c2 | tail -2
echo "== 1cta 256"; NVTE_B200_GEMM_1CTA=1 NVTE_B200_GEMM_1CTA_BN=256 bash scripts/gpu_kernel_bench.sh --only gemm_fp8_c2 | tail -2
echo "== 1cta 128"; NVTE_B200_GEMM_1CTA=1 NVTE_B200_GEMM_1CTA_BN=128 bash scripts/gpu_kernel_bench.sh --only gemm_fp8_c2 | tail -2
