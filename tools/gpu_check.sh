#!/bin/bash
# Run the GPU parity tests one test function per process (a device trap in one kernel then cannot take
# the others down), logs under gpurun_out/.  Usage: tools/gpu_check.sh [extra pytest args]
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
: > gpurun_out/gpu_check.log
ids=$(python -m pytest tests/test_gpu_kernels.py tests/test_gpu_modules.py -m gpu --collect-only -q 2>/dev/null | grep "::" | sed 's/\[.*//' | sort -u)
fail=0
for id in $ids; do
  echo "=== $id" >> gpurun_out/gpu_check.log
  timeout 900 python -m pytest "$id" -m gpu -q -x --no-header -p no:cacheprovider "$@" >> gpurun_out/gpu_check.log 2>&1
  rc=$?
  echo "--- rc=$rc $id" >> gpurun_out/gpu_check.log
  [ $rc -ne 0 ] && fail=1
done
grep -E "^--- rc=|passed|failed|Error|error|assert" gpurun_out/gpu_check.log | tail -80
exit $fail
