#!/bin/bash
# round 2, GPU call R (2 GPUs): multi-GPU tests and the driver's N = 2 bench command on the final tree
mkdir -p gpurun_out
python -m pytest tests/test_gpu_distributed.py -m gpu -q > gpurun_out/r_tests_dist.log 2>&1; echo "dist tests rc=$?"
tail -5 gpurun_out/r_tests_dist.log | cut -c1-200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r_bench2.json 2> gpurun_out/r_bench2.err
echo "bench rc=$?"; grep -E "Error|error" gpurun_out/r_bench2.err | head -5 | cut -c1-300
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r_bench2.json').read().strip().splitlines()[-1])
    print('ms/step',d['ms_per_step'],'e2e',d['e2e']['ms_per_step'],'config3',d['config3']['ms_per_step'])
    print('parity',json.dumps(d.get('parity'))[:600])
    print('spatial',{k:(round(v['ms_h'],2),round(v['ms_1gpu'],2)) for k,v in d['spatial']['blocks'].items()})
except Exception as e: print('failed',e)
PY
