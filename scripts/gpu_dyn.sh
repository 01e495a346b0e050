#!/bin/bash
# dynamic tile queue: parity with the queue forced on for every FP8 pair GEMM, then its effect on plain GEMMs
mkdir -p gpurun_out
echo "== FP8 GEMM tests, NVTE_B200_GEMM_DYNAMIC=1"
NVTE_B200_GEMM_DYNAMIC=1 timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_gemm_fullsize_gpu.py tests/test_linear_gpu.py -x -q -m gpu -k "fp8 or current or delayed or Float8" 2>&1 | tail -3 | cut -c1-300
echo "== norm 32x32"
timeout 300 python -m pytest tests/test_norm_gpu.py -x -q -m gpu -k "small_32x32" 2>&1 | tail -3 | cut -c1-300
echo "== static"; bash scripts/gpu_kernel_bench.sh --only gemm_fp8 | tail -4
echo "== dynamic"; NVTE_B200_GEMM_DYNAMIC=1 bash scripts/gpu_kernel_bench.sh --only gemm_fp8 | tail -4
