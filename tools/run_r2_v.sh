#!/bin/bash
# round 2, GPU call V (1 GPU): nlon = 1440 with all 721 modes on the tiled DFT kernels (three bands) - kernel tests, the
# spectral-loss tests and timing, then the whole GPU suite and smoke
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "dft" > gpurun_out/v_dft.log 2>&1; echo "dft rc=$?" | tee -a gpurun_out/v_dft.log
tail -30 gpurun_out/v_dft.log | cut -c1-220
timeout 300 python -m pytest tests/test_gpu_losses.py -m gpu -q > gpurun_out/v_losses.log 2>&1; echo "losses rc=$?" | tee -a gpurun_out/v_losses.log
tail -12 gpurun_out/v_losses.log | cut -c1-220
timeout 240 python tools/bench_h1_loss.py > gpurun_out/v_h1_bench.txt 2>&1; echo "bench_h1 rc=$?"; grep h1_loss gpurun_out/v_h1_bench.txt | cut -c1-250
timeout 400 python -m pytest tests -m gpu -q --deselect tests/test_gpu_losses.py > gpurun_out/v_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/v_tests.log
tail -4 gpurun_out/v_tests.log | cut -c1-200
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/v_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/v_smoke.log
