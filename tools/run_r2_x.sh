#!/bin/bash
# round 2, GPU call X (1 GPU): final single-GPU evidence - whole GPU suite, smoke, default bench, reference arm, launch list
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/x_no_alloc_alone.log 2>&1; echo "no_alloc alone rc=$?" | tee -a gpurun_out/x_no_alloc_alone.log
python -m pytest tests -m gpu -q > gpurun_out/x_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/x_tests.log
tail -6 gpurun_out/x_tests.log | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/x_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/x_smoke.log
tail -2 gpurun_out/x_smoke.log
python bench.py > gpurun_out/x_bench.json 2> gpurun_out/x_bench.err; rc=$?; echo "bench rc=$rc"
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/x_reference.json 2> gpurun_out/x_reference.err; echo "reference rc=$?"
cut -c1-200 gpurun_out/x_reference.json
if [ $rc -eq 0 ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1600 --csv --log-file gpurun_out/x_launches_bench.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-gpu-baseline --no-rollout --no-config3 --no-full-model --no-packed-master > gpurun_out/x_ncu.log 2>&1
  echo "ncu rc=$?"; wc -l gpurun_out/x_launches_bench.csv
fi
