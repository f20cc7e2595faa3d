#!/bin/bash
# round 2, GPU call H (8 GPUs): the driver's N = 8 command - data parallel step + parity / spatial as 2 DP x 4 h
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/h_bench$N.json 2> gpurun_out/h_bench$N.err
echo "bench rc=$?"
grep -E "Error|error|Traceback" gpurun_out/h_bench$N.err | head -10 | cut -c1-300
cut -c1-400 gpurun_out/h_bench$N.json
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/h_bench$N.json').read().strip().splitlines()[-1])
    print('ms/step',d['ms_per_step'],'e2e',d['e2e'])
    print('parity',json.dumps(d.get('parity'))[:1500])
    print('spatial',json.dumps(d.get('spatial'))[:2500])
    print('config3',d.get('config3'))
except Exception as e: print('failed',e)
PY
