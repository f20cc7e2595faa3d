#!/bin/bash
# round 2, GPU call E (4 GPUs): all distributed GPU tests (2- and 4-rank, CUDA-graph replay over peer memory), bench --gpus 4
mkdir -p gpurun_out
N=${1:-4}
python -m pytest tests/test_gpu_distributed.py -m gpu -q > gpurun_out/e_tests_dist.log 2>&1; tail -12 gpurun_out/e_tests_dist.log | cut -c1-300
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/e_bench$N.json 2> gpurun_out/e_bench$N.err
echo "bench rc=$?"; tail -4 gpurun_out/e_bench$N.err | cut -c1-300
python - <<PY
import json
d=json.loads(open('gpurun_out/e_bench$N.json').read().strip().splitlines()[-1])
print('ms/step',round(d['ms_per_step'],3),'e2e',round(d['e2e']['ms_per_step'],3),'cfg3',d.get('config3'))
print('parity',d.get('parity'))
s=d.get('spatial')
if s:
    for k,v in s['blocks'].items(): print(k,{a:round(b,3) for a,b in v.items()})
    print(s['lat_to_chan_721'])
PY
NCCL_DEBUG=INFO python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 bench.py --gpus $N --steps 3 --warmup 3 --no-spatial --no-config3 > gpurun_out/e_bench${N}_nccldebug.json 2> gpurun_out/e_nccl_debug.log
grep -i -E "nvls|algo|channels" gpurun_out/e_nccl_debug.log | head -12 | cut -c1-200
grep -c . gpurun_out/e_nccl_debug.log; head -c 3000000 gpurun_out/e_nccl_debug.log > gpurun_out/e_nccl_debug_head.log; rm gpurun_out/e_nccl_debug.log
