#!/bin/bash
# A/B of the store path of csrc/pointwise.cu on the GPU box: segment boundaries aligned to 128 bytes (SFNO_PW_ALIGN) and
# streaming stores (SFNO_PW_STCS), with the fast erf.  -> gpurun_out/ab_pw_store.txt
mkdir -p gpurun_out
: > gpurun_out/ab_pw_store.txt
for v in "0 0" "1 0" "0 1" "1 1"; do
  set -- $v
  rm -f makani_b200/csrc/pointwise.o
  make -C makani_b200/csrc EXTRA="-DSFNO_FAST_ERF=1 -DSFNO_PW_ALIGN=$1 -DSFNO_PW_STCS=$2" > /dev/null 2>&1 || { echo "build failed ($v)"; exit 1; }
  { echo "=== ALIGN=$1 STCS=$2"; timeout 60 python tools/bench_pointwise_kernels.py; } >> gpurun_out/ab_pw_store.txt 2>&1
done
grep -E "===|721x1440" gpurun_out/ab_pw_store.txt
rm -f makani_b200/csrc/pointwise.o
make -C makani_b200/csrc > /dev/null 2>&1
