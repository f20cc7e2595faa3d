#!/bin/bash
# round 2: N = 8 data-parallel step with / without binding every rank to the CPUs next to its GPU (e2e is host-bound there)
mkdir -p gpurun_out
lscpu | grep -E "^CPU\(s\)|Socket|NUMA" | head -8 > gpurun_out/numa8_lscpu.txt; cat gpurun_out/numa8_lscpu.txt
for v in 1 0; do
  MAKANI_B200_BENCH_NUMA=$v timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29595 bench.py --gpus 8 --steps 10 --warmup 3 --no-spatial --no-config3 > gpurun_out/numa8_$v.json 2> gpurun_out/numa8_$v.err
  echo "numa=$v rc=$?"; grep -E "Error|error" gpurun_out/numa8_$v.err | head -3 | cut -c1-300
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/numa8_$v.json').read().strip().splitlines()[-1])
    print('numa=$v ms/step',round(d['ms_per_step'],3),'e2e',round(d['e2e']['ms_per_step'],2),'|',d['config'].get('host_affinity'))
except Exception as e: print('failed',e)
PY
done
