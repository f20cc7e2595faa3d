#!/bin/bash
# round 2, last GPU call (1 GPU): final tree - whole GPU suite, no-alloc test alone, smoke, default bench
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -q > gpurun_out/f_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/f_tests.log
tail -3 gpurun_out/f_tests.log | cut -c1-200
timeout 100 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/f_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?" | tee -a gpurun_out/f_no_alloc_alone.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/f_smoke.log
timeout 300 python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench rc=$?"
cut -c1-330 gpurun_out/f_bench.json
