#!/bin/bash
# ncu evidence for one round: launch lists (share of step) + one full capture per kernel family.
# Usage (under gpurun): bash tools/ncu_round.sh <tag>
tag=${1:-r1}
mkdir -p gpurun_out
for which in internal block0 block7; do
  python tools/prof_block.py $which 2 > gpurun_out/plain_$which.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_launches_$which.csv \
      python tools/prof_block.py $which 2 > gpurun_out/ncu_l_$which.log 2>&1
done
python tools/prof_block.py internal 1 > gpurun_out/plain_i1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -c 13 -o gpurun_out/${tag}_full_internal -f \
    python tools/prof_block.py internal 1 > gpurun_out/ncu_f_internal.log 2>&1
python tools/prof_block.py block0 1 > gpurun_out/plain_b1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"dft_|gemm_tc_kernel" -c 10 -o gpurun_out/${tag}_full_block0 -f \
    python tools/prof_block.py block0 1 > gpurun_out/ncu_f_block0.log 2>&1
python tools/prof_block.py block7 1 > gpurun_out/plain_b7.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"dft_inv_p2_kernel|gemm_tc_kernel" -c 6 -o gpurun_out/${tag}_full_block7 -f \
    python tools/prof_block.py block7 1 > gpurun_out/ncu_f_block7.log 2>&1
ls -la gpurun_out
