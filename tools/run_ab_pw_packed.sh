#!/bin/bash
# GPU check of the packed (FFMA2) GELU in csrc/pointwise.cu: parity tests + per-kernel timings.  -> gpurun_out/pw_packed.txt
mkdir -p gpurun_out
{
  timeout 300 python -m pytest tests/test_gpu_pointwise.py tests/test_gpu_pointwise_mlp.py tests/test_gpu_z_block.py -m gpu -q 2>&1 | tail -3
  timeout 60 python tools/bench_pointwise_kernels.py
} > gpurun_out/pw_packed.txt 2>&1
cat gpurun_out/pw_packed.txt
