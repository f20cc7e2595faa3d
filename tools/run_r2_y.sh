#!/bin/bash
# round 2, GPU call Y (1 GPU): register twiddles at R = 1 by default - whole GPU suite, DFT kernel timings, smoke, default bench
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q > gpurun_out/y_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/y_tests.log
tail -4 gpurun_out/y_tests.log | cut -c1-200
timeout 120 python tools/bench_dft.py > gpurun_out/y_dft.txt 2>&1; cat gpurun_out/y_dft.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/y_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/y_smoke.log
timeout 300 python bench.py > gpurun_out/y_bench.json 2> gpurun_out/y_bench.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/y_bench.json
