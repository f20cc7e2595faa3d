#!/bin/bash
# round 2, GPU call Z (2 GPUs): peer-memory gradient exchange (gather unpack) parity against NCCL + the default N = 2 bench
mkdir -p gpurun_out
run() { tag=$1; shift
  env "$@" timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29591 bench.py --gpus 2 --steps 10 --warmup 3 --no-spatial --no-config3 > gpurun_out/z_bench2_$tag.json 2> gpurun_out/z_bench2_$tag.err
  echo "$tag rc=$?"; grep -E "Error|error" gpurun_out/z_bench2_$tag.err | head -3 | cut -c1-300
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/z_bench2_$tag.json').read().strip().splitlines()[-1])
    print('$tag ms/step',round(d['ms_per_step'],3),'parity',d.get('parity'))
except Exception as e: print('$tag failed',e)
PY
}
run peer MAKANI_B200_DP_EXCHANGE=peer
run nccl MAKANI_B200_DP_EXCHANGE=nccl
