#!/bin/bash
# round 2, GPU call W (1 GPU): spectral loss (GeometricH1Loss) parity tests + timing, then the whole GPU suite and smoke
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_losses.py -m gpu -q > gpurun_out/w_losses.log 2>&1; echo "losses rc=$?" | tee -a gpurun_out/w_losses.log
tail -25 gpurun_out/w_losses.log | cut -c1-220
timeout 240 python tools/bench_h1_loss.py > gpurun_out/w_h1_bench.txt 2>&1; echo "bench_h1 rc=$?"; cat gpurun_out/w_h1_bench.txt | cut -c1-250
timeout 400 python -m pytest tests -m gpu -q --deselect tests/test_gpu_losses.py > gpurun_out/w_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/w_tests.log
tail -4 gpurun_out/w_tests.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/w_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/w_smoke.log
