#!/bin/bash
# round 2, GPU call N (1 GPU): full ncu capture of the five kernels of one internal block forward (GEMM memory-path metrics)
mkdir -p gpurun_out
python tools/prof_block.py internal 2 fwd > gpurun_out/n_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/n_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_kernel|dft_" --launch-skip 5 -c 5 -o gpurun_out/n_internal_fwd -f \
    python tools/prof_block.py internal 2 fwd > gpurun_out/n_ncu.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/n_internal_fwd.ncu-rep
