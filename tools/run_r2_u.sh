#!/bin/bash
# round 2, GPU call U (1 GPU): launch list of the spectral loss at 721x1440 (fused leg, 73 channels, forward + backward)
mkdir -p gpurun_out
timeout 80 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/u_h1_launches.csv \
    python tools/bench_h1_loss.py 1 721 fused > gpurun_out/u_ncu.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/u_h1_launches.csv; tail -2 gpurun_out/u_ncu.log | cut -c1-200
