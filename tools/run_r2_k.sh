#!/bin/bash
# round 2, GPU call K (1 GPU): pointwise-MLP parity tests, whole-model step in three variants, kernel table of the fused one
mkdir -p gpurun_out
python -m pytest tests/test_gpu_pointwise_mlp.py tests/test_gpu_pointwise.py -m gpu -q -x > gpurun_out/k_tests.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/k_tests.log | cut -c1-250
python - > gpurun_out/k_full_model.json 2> gpurun_out/k_full_model.err <<PY
import sys, json, torch
sys.path.insert(0, 'tools')
from benchlib import sfno_model
print(json.dumps(sfno_model.measure(torch.device('cuda:0'))))
PY
echo "full model rc=$?"; cat gpurun_out/k_full_model.json | cut -c1-1500; tail -5 gpurun_out/k_full_model.err | cut -c1-300
python tools/prof_full_model.py fused > gpurun_out/k_prof_full_fused.txt 2>&1; head -45 gpurun_out/k_prof_full_fused.txt | cut -c1-200
