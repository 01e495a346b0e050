#!/bin/bash
mkdir -p gpurun_out
BIN=oracle/_ref/ref_operator_tests
for f in "OperatorTest/CTTestSuite.*" "OperatorTest/DequantizeMXFP8TestSuite.*" "OperatorTest/DequantizeNVFP4TestSuite.*" "OperatorTest/NormTestSuite.*" "OperatorTest/MxNormTestSuite.*"; do
  n=$(echo $f | tr '/.*' '___')
  timeout 600 $BIN --gtest_filter="$f" --gtest_brief=1 2>&1 | grep -v "^\[  SKIPPED\|^\[       OK\|^\[ RUN" | head -400 > gpurun_out/gt_$n.log
done
