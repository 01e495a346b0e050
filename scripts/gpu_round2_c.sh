#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_norm_gpu.py tests/test_nvfp4_gpu.py tests/test_quantize_gpu.py tests/test_guard_gpu.py tests/test_gemm_gpu.py -x -q -m gpu 2>&1 | tail -4 | cut -c1-300
bash scripts/gpu_kernel_bench.sh --only nvfp4_c3 | tail -6
SUITE_TIMEOUT=300 bash scripts/gpu_ref_gtests.sh > /dev/null 2>&1; cat gpurun_out/r02_ref_gtests.txt
