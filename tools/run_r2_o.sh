#!/bin/bash
# round 2, GPU call O (1 GPU): packed-master tests + bench line with the packed_master and full_model keys
mkdir -p gpurun_out
python -m pytest tests/test_gpu_modules.py tests/test_gpu_pointwise_mlp.py -m gpu -q -x -k "packed_master or pointwise or fused_model" > gpurun_out/o_tests.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/o_tests.log | cut -c1-250
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/o_bench.json 2> gpurun_out/o_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/o_bench.err | cut -c1-300
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/o_bench.json').read().strip().splitlines()[-1])
    print('ms/step',d['ms_per_step'],'packed',d['packed_master'] and (d['packed_master']['ms_per_step'], d['packed_master']['launches_per_step']))
    fm=d['full_model']; print({k:round(fm[k]['ms_per_step'],2) for k in ('torch_glue','fused_norm','fused')} if 'fused' in fm else fm)
    print('wall', d['config'].get('wall_s'))
except Exception as e: print('failed',e)
PY
