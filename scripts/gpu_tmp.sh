timeout 120 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 120 python -m pytest tests/test_nvfp4_gpu.py -m gpu -x -q -k gemm 2>&1 | tail -3
timeout 200 python scripts/cublaslt_ref.py 2>&1 | tail -5
NVTE_B200_GEMM_NO_MULTICAST=1 timeout 200 python scripts/cublaslt_ref.py 2>&1 | tail -5
