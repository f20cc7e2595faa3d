#!/bin/bash
# round 2, GPU call I (1 GPU): bench with the whole-model key, CTA-order A/B of the weight layout kernels, compute-sanitizer
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/i_bench.json').read().strip().splitlines()[-1])
    print('ms/step',d['ms_per_step'],'rollout',d['rollout']['ms_per_step_eager'],d['rollout']['ms_per_step_graph'])
    print('full_model',json.dumps(d.get('full_model'))[:1800])
except Exception as e: print('failed',e)
PY
tail -3 gpurun_out/i_bench.err | cut -c1-300
for o in 0 1 0 1; do SFNO_WLAYOUT_ORDER=$o python tools/bench_weight_layout.py 2>&1 | tail -1 | tee -a gpurun_out/i_wlayout.txt; done
bash tools/sanitize.sh
