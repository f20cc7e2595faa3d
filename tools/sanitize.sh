#!/bin/bash
# compute-sanitizer passes over the small GPU parity tests (the reference has no sanitizer runs at all).
#   gpurun --timeout 900 -- 'bash tools/sanitize.sh'     -> gpurun_out/sanitize_{memcheck,racecheck,synccheck}.log
# NOTE (round 2): compute-sanitizer is closed on the B200 pool this repository is measured on (the wrapper on the box
# prints that runs under it have left GPUs needing a reset), so this script produced no log there.
# Small cases only: the tools slow kernels down 10-100x.
mkdir -p gpurun_out
SEL='test_real_sht_forward_backward and 32-64 or test_spectral_conv_matches_reference_module_fixtures or test_fused_instance_norm_forward_backward and 12-24 or test_static_variant_equals'
for tool in memcheck racecheck synccheck; do
  timeout 280 compute-sanitizer --tool $tool --error-exitcode 9 \
    python -m pytest tests/test_gpu_modules.py tests/test_gpu_pointwise.py tests/test_zz_no_alloc.py -m gpu -q -x -k "$SEL" \
    > gpurun_out/sanitize_$tool.log 2>&1
  echo "$tool rc=$?  $(grep -E 'ERROR SUMMARY|passed|failed' gpurun_out/sanitize_$tool.log | tail -2 | tr '\n' ' ')"
done
