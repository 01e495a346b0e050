#!/bin/bash
# Same-box comparison: the reference's own sm_100a kernels (baseline/_ref/libte_refkernels.so; its GEMM entry is
# cuBLASLt) and ours, through the same harness and the same nvte_* calls, every case timed INTERLEAVED in one process.
# Output: gpurun_out/r02_kernel_bench.jsonl + a table. Extra arguments go to kernel_bench (e.g. --only gemm_).
mkdir -p gpurun_out
timeout 1200 baseline/_ref/kernel_bench baseline/_ref/libte_refkernels.so=reference transformerengine_b200/libtransformer_engine.so=ours "$@" \
  > gpurun_out/r02_kernel_bench.jsonl 2> gpurun_out/r02_kernel_bench.err
tail -3 gpurun_out/r02_kernel_bench.err
python scripts/kernel_bench_table.py gpurun_out/r02_kernel_bench.jsonl gpurun_out/r02_kernel_bench_table.json | tee gpurun_out/r02_kernel_bench_table.txt
