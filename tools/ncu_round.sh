#!/bin/bash
# ncu evidence for one round: launch lists (share of step) + one full capture per kernel family.
# Usage (under gpurun): bash tools/ncu_round.sh <tag>
# The summaries are made on the box (gpurun_out/ is limited to 64 MiB): text tables + the traffic table land in
# gpurun_out/profiles_<tag>/; only the report with the dominant kernel (block7: 721x1440 inverse DFT, Legendre
# synthesis) is kept as .ncu-rep for source-level reading.
tag=${1:-r1}
out=gpurun_out/profiles_$tag
mkdir -p $out
for which in internal block0 block7; do
  python tools/prof_block.py $which 2 > gpurun_out/plain_$which.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $out/${tag}_launches_$which.csv \
      python tools/prof_block.py $which 2 > gpurun_out/ncu_l_$which.log 2>&1
done
rm -f profiles/ncu_traffic.json
python tools/prof_block.py internal 1 > gpurun_out/plain_i1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -c 13 -o /tmp/${tag}_full_internal -f \
    python tools/prof_block.py internal 1 > gpurun_out/ncu_f_internal.log 2>&1
python tools/prof_block.py block0 1 > gpurun_out/plain_b1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"dft_|gemm_tc_kernel" -c 10 -o /tmp/${tag}_full_block0 -f \
    python tools/prof_block.py block0 1 > gpurun_out/ncu_f_block0.log 2>&1
python tools/prof_block.py block7 1 > gpurun_out/plain_b7.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"dft_inv_p2_kernel|gemm_tc_kernel" -c 6 -o /tmp/${tag}_full_block7 -f \
    python tools/prof_block.py block7 1 > gpurun_out/ncu_f_block7.log 2>&1
for which in internal block0 block7; do
  python tools/ncu_summary.py /tmp/${tag}_full_$which.ncu-rep > $out/${tag}_ncu_full_$which.txt 2>&1
done
python tools/ncu_traffic.py /tmp/${tag}_full_internal.ncu-rep /tmp/${tag}_full_block0.ncu-rep /tmp/${tag}_full_block7.ncu-rep > $out/ncu_traffic.txt 2>&1
cp profiles/ncu_traffic.json $out/ncu_traffic.json
cp /tmp/${tag}_full_block7.ncu-rep gpurun_out/
du -sh gpurun_out; ls -la $out
