#!/bin/bash
# Drop-in proof (SURVEY 7.1 step 2 / VERDICT r1 next-3): the reference's OWN C++ operator tests
# (tests/cpp/operator/*.cu, built by `make -C oracle tests` against transformerengine_b200/libtransformer_engine.so)
# run on the B200, one gtest suite at a time. Writes gpurun_out/r02_ref_gtests.txt (pass / fail counts per suite)
# and the failing test names under gpurun_out/r02_ref_gtests_failures.txt.
mkdir -p gpurun_out
BIN=oracle/_ref/ref_operator_tests
OUT=gpurun_out/r02_ref_gtests.txt; FAIL=gpurun_out/r02_ref_gtests_failures.txt
: > $OUT; : > $FAIL
printf "%-72s %7s %7s %7s %8s\n" "suite (reference tests/cpp/operator)" passed failed skipped seconds >> $OUT
for suite in $($BIN --gtest_list_tests | grep -v "^  " | grep -v "^Running" | grep -v "${SKIP_SUITES:-4Over6}" | grep -E "${ONLY_SUITES:-.}"); do
  t0=$(date +%s)
  timeout ${SUITE_TIMEOUT:-900} $BIN --gtest_filter="${suite}*" --gtest_brief=1 > /tmp/gt.log 2>&1
  rc=$?
  t1=$(date +%s)
  p=$(grep -oE "\[  PASSED  \] [0-9]+" /tmp/gt.log | grep -oE "[0-9]+$" | tail -1)
  f=$(grep -oE "^ *[0-9]+ FAILED TEST" /tmp/gt.log | grep -oE "[0-9]+" | tail -1)
  s=$(grep -oE "\[  SKIPPED \] [0-9]+ test" /tmp/gt.log | grep -oE "[0-9]+" | tail -1)
  [ $rc -ne 0 ] && [ -z "$f" ] && f="rc=$rc"
  printf "%-72s %7s %7s %7s %8s\n" "$suite" "${p:-0}" "${f:-0}" "${s:-0}" "$((t1 - t0))" >> $OUT
  grep -E "^\[  FAILED  \] .*GetParam|^\[  FAILED  \] [A-Za-z]" /tmp/gt.log | sort -u | head -40 >> $FAIL
  if [ -n "$f" ] && [ "$f" != "0" ]; then grep -m6 -B1 -A6 "Failure\|exception\|error" /tmp/gt.log | head -60 >> $FAIL; fi
done
cat $OUT
echo "---- first failures"; head -60 $FAIL
