#!/bin/bash
# Round-2 first GPU pass: new parity tests, the comparator harness on both libraries, then the whole GPU suite.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r02_smi.txt
timeout 1500 python -m pytest tests/test_gemm_fullsize_gpu.py tests/test_cabi.py -x -q -m gpu --durations=20 > gpurun_out/r02_pytest_fullsize.log 2>&1
tail -5 gpurun_out/r02_pytest_fullsize.log
timeout 900 python -m pytest tests/test_nvfp4_gpu.py tests/test_linear_gpu.py tests/test_quantize_gpu.py -x -q -m gpu > gpurun_out/r02_pytest_new.log 2>&1
tail -5 gpurun_out/r02_pytest_new.log
timeout 600 baseline/_ref/kernel_bench baseline/_ref/libte_refkernels.so --name reference > gpurun_out/r02_kernel_bench_reference.jsonl 2> gpurun_out/r02_kernel_bench_reference.err
tail -3 gpurun_out/r02_kernel_bench_reference.err
timeout 600 baseline/_ref/kernel_bench transformerengine_b200/libtransformer_engine.so --name ours > gpurun_out/r02_kernel_bench_ours.jsonl 2> gpurun_out/r02_kernel_bench_ours.err
tail -3 gpurun_out/r02_kernel_bench_ours.err
cat gpurun_out/r02_kernel_bench_reference.jsonl gpurun_out/r02_kernel_bench_ours.jsonl | cut -c1-250
