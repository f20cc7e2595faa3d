#!/bin/bash
# round 2: data-parallel step at N = 8 with the peer-memory gradient exchange (parity against NCCL inside the bench)
mkdir -p gpurun_out
NCCL_DEBUG=WARN MAKANI_B200_DP_EXCHANGE=peer timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29593 bench.py --gpus 8 --steps 10 --warmup 3 --no-spatial --no-config3 > gpurun_out/dp8_peer.json 2> gpurun_out/dp8_peer.err
echo "rc=$?"; grep -E "Error|error" gpurun_out/dp8_peer.err | head -3 | cut -c1-300
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/dp8_peer.json').read().strip().splitlines()[-1])
    print('peer N=8 ms/step',round(d['ms_per_step'],3),'e2e',round(d['e2e']['ms_per_step'],2),'parity',d.get('parity'), d['config'].get('dp_gradient_exchange','')[:80])
except Exception as e: print('failed',e)
PY
