#!/bin/bash
# A/B of the packed fp32 complex arithmetic (FADD2 / FFMA2) and of twiddles formed in registers in the DFT kernels.  -> gpurun_out/ab_f32x2.txt
mkdir -p gpurun_out
: > gpurun_out/ab_f32x2.txt
for v in "0 0" "1 0" "1 1"; do
  set -- $v
  rm -f makani_b200/csrc/dft.o
  rm -f makani_b200/csrc/dft.o; make -C makani_b200/csrc EXTRA="-DSFNO_F32X2=$1 -DSFNO_TW_REGS=$2" > /dev/null 2>&1 || { echo "build failed ($v)"; exit 1; }
  {
    echo "=== SFNO_F32X2=$1 SFNO_TW_REGS=$2"
    timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_modules.py tests/test_gpu_known_answer.py -m gpu -q -x 2>&1 | tail -2
    timeout 120 python tools/bench_dft.py
  } >> gpurun_out/ab_f32x2.txt 2>&1
done
cat gpurun_out/ab_f32x2.txt
rm -f makani_b200/csrc/dft.o
make -C makani_b200/csrc > /dev/null 2>&1
