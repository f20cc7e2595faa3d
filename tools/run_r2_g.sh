#!/bin/bash
# round 2, GPU call G (1 GPU): the allocation-free entry point alone in a fresh process, the whole GPU suite, smoke,
# the default bench, the reference arm, and the ncu launch list of the same bench command
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/g_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?" | tee -a gpurun_out/g_no_alloc_alone.log
tail -5 gpurun_out/g_no_alloc_alone.log
python -m pytest tests -m gpu -q > gpurun_out/g_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/g_tests.log
tail -8 gpurun_out/g_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/g_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/g_smoke.log
tail -3 gpurun_out/g_smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/g_bench.json 2> gpurun_out/g_bench.err; rc=$?; echo "bench rc=$rc"
cut -c1-600 gpurun_out/g_bench.json
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/g_reference.json 2> gpurun_out/g_reference.err; echo "reference rc=$?"
cut -c1-300 gpurun_out/g_reference.json
if [ $rc -eq 0 ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1600 --csv --log-file gpurun_out/g_launches_bench.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-gpu-baseline --no-rollout --no-config3 --no-full-model > gpurun_out/g_ncu.log 2>&1
  echo "ncu rc=$?"; wc -l gpurun_out/g_launches_bench.csv
fi
