#!/bin/bash
# compute-sanitizer memcheck + racecheck over scripts/sanitizer_target.py; summaries -> gpurun_out/r02_sanitizer_*.txt
mkdir -p gpurun_out
python scripts/sanitizer_target.py > gpurun_out/r02_sanitizer_plain.txt 2>&1 || { tail -20 gpurun_out/r02_sanitizer_plain.txt; exit 1; }
for tool in memcheck racecheck; do
  for part in mxfp8 nvfp4 norm gemm; do
    echo "=== $tool $part" >> gpurun_out/r02_sanitizer_$tool.txt
    timeout 900 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitizer_target.py $part 2>&1 | grep -v "^$" | tail -25 >> gpurun_out/r02_sanitizer_$tool.txt
  done
  grep -E "===|ERROR SUMMARY|RACECHECK SUMMARY|^ok|hazard|Invalid|Error" gpurun_out/r02_sanitizer_$tool.txt | head -60
done
