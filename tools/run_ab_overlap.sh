#!/bin/bash
# A/B of the weight-path overlap on the GPU box:
#   base                                   : 64 x 64 pack / unpack tiles, everything on one stream
#   SFNO_WPACK_TL=32                       : 16 KB tiles only (do the smaller tiles cost anything by themselves?)
#   + MAKANI_B200_OVERLAP_WEIGHTS=1        : unpack on a side stream during the rest of backward, re-packs prefetched
# Parity tests with the variant on, then the device-resident bench.   gpurun --timeout 400 -- 'bash tools/run_ab_overlap.sh'
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" timeout 120 python -m pytest tests/test_gpu_modules.py tests/test_gpu_kernels.py -m gpu -q -x > gpurun_out/ab_overlap_${tag}_tests.log 2>&1
  echo "$tag tests rc=$? $(tail -1 gpurun_out/ab_overlap_${tag}_tests.log)"
  env "$@" timeout 90 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/ab_overlap_${tag}.json 2> gpurun_out/ab_overlap_${tag}.err
  echo "$tag bench rc=$? $(python tools/show_bench.py gpurun_out/ab_overlap_${tag}.json 2>&1 | head -1)"
}
run base X=1
run tl32 SFNO_WPACK_TL=32
run overlap SFNO_WPACK_TL=32 MAKANI_B200_OVERLAP_WEIGHTS=1
