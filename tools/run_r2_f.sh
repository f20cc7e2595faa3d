#!/bin/bash
# round 2, GPU call F (N GPUs): data-parallel step with the peer-memory gradient exchange vs the NCCL all-reduce
mkdir -p gpurun_out
N=${1:-2}
run() { tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29566 bench.py --gpus $N --steps 10 --warmup 3 --no-spatial --no-config3 > gpurun_out/f_bench${N}_$tag.json 2> gpurun_out/f_bench${N}_$tag.err
  echo "$tag rc=$?"; grep -E "Error|error" gpurun_out/f_bench${N}_$tag.err | head -5 | cut -c1-300
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/f_bench${N}_$tag.json').read().strip().splitlines()[-1])
    print('$tag ms/step',round(d['ms_per_step'],3),'e2e',round(d['e2e']['ms_per_step'],3),'parity',d.get('parity'))
    for k in d['kernels'][:16]: print('   ',k['kernel'],k['shape'].get('K'),k['ms_per_step'])
except Exception as e: print('$tag failed',e)
PY
}
run direct MAKANI_B200_DP_EXCHANGE=peer

run nccl MAKANI_B200_DP_EXCHANGE=nccl
