#!/bin/bash
# Same-box comparison: the reference's own sm_100a kernels (baseline/_ref/libte_refkernels.so) and ours through
# the same harness and the same nvte_* calls. Output: gpurun_out/r02_kernel_bench_{reference,ours}.jsonl
mkdir -p gpurun_out
for lib in reference ours; do
  so=baseline/_ref/libte_refkernels.so; [ $lib = ours ] && so=transformerengine_b200/libtransformer_engine.so
  timeout 900 baseline/_ref/kernel_bench $so --name $lib "$@" > gpurun_out/r02_kernel_bench_$lib.jsonl 2> gpurun_out/r02_kernel_bench_$lib.err
  tail -3 gpurun_out/r02_kernel_bench_$lib.err
done
python scripts/kernel_bench_table.py gpurun_out/r02_kernel_bench_reference.jsonl gpurun_out/r02_kernel_bench_ours.jsonl | tee gpurun_out/r02_kernel_bench_table.txt
