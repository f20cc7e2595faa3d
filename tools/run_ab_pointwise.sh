#!/bin/bash
# A/B of the GELU arithmetic in csrc/pointwise.cu on the GPU box: default (erff) vs -DSFNO_FAST_ERF=1 (A&S 7.1.26).
# For each variant: rebuild pointwise.o + the .so, run the pointwise / block parity tests, then the per-kernel timings.
#   gpurun --timeout 400 -- 'bash tools/run_ab_pointwise.sh'      -> gpurun_out/ab_pointwise_{0,1}.log
mkdir -p gpurun_out
for v in 0 1; do
  rm -f makani_b200/csrc/pointwise.o
  make -C makani_b200/csrc EXTRA=-DSFNO_FAST_ERF=$v > /dev/null 2>&1 || { echo "build failed (variant $v)"; exit 1; }
  {
    echo "=== SFNO_FAST_ERF=$v"
    timeout 200 python -m pytest tests/test_gpu_pointwise.py tests/test_gpu_pointwise_mlp.py tests/test_gpu_z_block.py -m gpu -q 2>&1 | tail -3
    timeout 60 python tools/bench_pointwise_kernels.py
  } > gpurun_out/ab_pointwise_$v.log 2>&1
  cat gpurun_out/ab_pointwise_$v.log
done
# leave the default build in place
rm -f makani_b200/csrc/pointwise.o
make -C makani_b200/csrc > /dev/null 2>&1
