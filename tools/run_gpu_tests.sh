#!/bin/bash
# Run the GPU test-suite file by file (a CUDA fault in one file must not hide the others); logs to gpurun_out/.
mkdir -p gpurun_out
: > gpurun_out/gpu_tests.log
for f in "$@"; do
  echo "=== $f" >> gpurun_out/gpu_tests.log
  timeout 900 python -m pytest "$f" -m gpu -q --timeout 800 -s --tb=short 2>&1 | grep -vE "^\s*$" | tail -400 >> gpurun_out/gpu_tests.log
done
grep -E "^===|passed|failed|error" gpurun_out/gpu_tests.log | tail -20
