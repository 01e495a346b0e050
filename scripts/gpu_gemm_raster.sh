#!/bin/bash
# dev aid: L2-raster sweep of the pair GEMM, each setting interleaved with cuBLASLt (the reference's GEMM entry)
mkdir -p gpurun_out; : > gpurun_out/r02_gemm_raster.txt
for g in ${RASTERS:-1 8 12 16}; do
  NVTE_B200_GEMM_RASTER=$g timeout 600 baseline/_ref/kernel_bench baseline/_ref/libte_refkernels.so=reference transformerengine_b200/libtransformer_engine.so=ours --only gemm_ > gpurun_out/raster_$g.jsonl 2>/dev/null
  echo "== raster $g" | tee -a gpurun_out/r02_gemm_raster.txt
  python scripts/kernel_bench_table.py gpurun_out/raster_$g.jsonl | tee -a gpurun_out/r02_gemm_raster.txt
done
