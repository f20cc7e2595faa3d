#!/bin/bash
# round 2, GPU call C (2 GPUs): fixed tests, 2-GPU distributed tests (incl. CUDA-graph replay over peer memory), bench --gpus 2
# with parity / spatial, NCCL CTA-cap A/B for the data-parallel gradient all-reduce
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_modules.py tests/test_gpu_contraction.py -m gpu -q -k "diagonal or weight_update or contract" > gpurun_out/c_tests_fixed.log 2>&1; tail -3 gpurun_out/c_tests_fixed.log
python -m pytest tests/test_gpu_distributed.py -m gpu -q > gpurun_out/c_tests_dist.log 2>&1; tail -8 gpurun_out/c_tests_dist.log
run() { tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 $EXTRA > gpurun_out/c_bench2_$tag.json 2> gpurun_out/c_bench2_$tag.err
  echo "$tag rc=$?"; python -c "
import json
try:
    d=json.load(open('gpurun_out/c_bench2_$tag.json')); print('$tag ms/step',round(d['ms_per_step'],3),'e2e',round(d['e2e']['ms_per_step'],3),'cfg3',d.get('config3'),'parity',d.get('parity'))
    s=d.get('spatial')
    if s: print({k:(round(v['ms_h'],3),round(v['ms_1gpu'],3)) for k,v in s['blocks'].items()}, s.get('lat_to_chan_721'))
except Exception as e: print('$tag failed', e)
"; tail -3 gpurun_out/c_bench2_$tag.err; }
EXTRA="" run full A=1
EXTRA="--no-spatial --no-config3" run ctas4 NCCL_MAX_CTAS=4
EXTRA="--no-spatial --no-config3" run ctas8 NCCL_MAX_CTAS=8
EXTRA="--no-spatial --no-config3" run ctas2 NCCL_MAX_CTAS=2
