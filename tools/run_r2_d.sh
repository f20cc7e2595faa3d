#!/bin/bash
# round 2, GPU call D (2 GPUs): bisect the multi-GPU worker failure
mkdir -p gpurun_out
W="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 tests/distributed/gpu_worker.py 2 1 37 72 equiangular 6"
MAKANI_TEST_SKIP_GRAPH=1 $W > gpurun_out/d_worker_nograph.log 2>&1; echo "nograph rc=$?"; grep -E "DIST_GPU|Error|error:|assert" gpurun_out/d_worker_nograph.log | head -8 | cut -c1-250
CUDA_LAUNCH_BLOCKING=1 $W > gpurun_out/d_worker_graph.log 2>&1; echo "graph rc=$?"; grep -E "DIST_GPU|Error|error:|assert|File \"/root|line " gpurun_out/d_worker_graph.log | head -24 | cut -c1-250
