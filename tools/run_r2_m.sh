#!/bin/bash
# round 2, GPU call M (1 GPU): full ncu capture (with source counters) of the R = 3 forward DFT and the R = 1 forward DFT
mkdir -p gpurun_out
python tools/prof_block.py block0 1 fwd > gpurun_out/m_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/m_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"dft_fwd_p2_kernel|dft_inv_p2_kernel" -c 2 -o gpurun_out/m_dft_block0 -f \
    python tools/prof_block.py block0 1 fwd > gpurun_out/m_ncu.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/m_dft_block0.ncu-rep
ncu -i gpurun_out/m_dft_block0.ncu-rep --page details --csv 2>/dev/null | grep -E "Issue Slots Busy|Executed Ipc|No Eligible|Eligible Warps|Active Warps|Stall|L1/TEX Hit|Registers|Achieved Occupancy|Duration|Bank conflict|Shared" | cut -c1-300 | head -60
