#!/bin/bash
# Runs every GPU test file in its own process (a faulting kernel then cannot take the other files down with it) and
# collects one summary line per file plus the failures' tracebacks under gpurun_out/.  Usage: tools/gpu_run_tests.sh [files...]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
FILES="$@"
if [ -z "$FILES" ]; then
  FILES="tests/test_gpu_parity_r2.py tests/test_gpu_kmeans_tc.py tests/test_gpu_kmeans.py tests/test_gpu_logreg.py tests/test_golden.py tests/test_gpu_guards.py tests/test_gpu_evaluation.py tests/test_gpu_multirank.py tests/test_zz_no_alloc.py"
fi
: > gpurun_out/tests_summary.txt
for f in $FILES; do
  name=$(basename $f .py)
  timeout 1500 python -m pytest $f -q -m gpu -s -rA --tb=short > gpurun_out/$name.log 2>&1
  rc=$?
  echo "$name rc=$rc $(tail -1 gpurun_out/$name.log)" >> gpurun_out/tests_summary.txt
  grep -E "^(FAILED|ERROR)|full-size" gpurun_out/$name.log >> gpurun_out/tests_summary.txt
done
cat gpurun_out/tests_summary.txt
