#!/bin/bash
# Everything that was queued when the round-1 GPU budget ran out, in one single-GPU call (~8 min of box time):
#   gpurun --timeout 900 -- 'bash tools/run_next_round_first.sh'
# Results under gpurun_out/next_*.  (Two-GPU items - the sharded fused norm in tests/distributed/gpu_worker.py -
# go through tools/run_dist2.sh.)
mkdir -p gpurun_out
timeout 120 python -m pytest tests -m gpu -q > gpurun_out/next_tests.log 2>&1; echo "gpu tests rc=$? $(tail -1 gpurun_out/next_tests.log)"
timeout 150 python tools/bench_torch_gpu.py --steps 5 --warmup 2 > gpurun_out/next_torch_gpu.json 2> gpurun_out/next_torch_gpu.err; echo "torch-gpu bar rc=$?"; cat gpurun_out/next_torch_gpu.json
timeout 60 python tools/bench_rollout.py 40 > gpurun_out/next_rollout.json 2> gpurun_out/next_rollout.err; echo "rollout rc=$?"; cat gpurun_out/next_rollout.json
timeout 120 bash tools/run_ab_pointwise.sh > gpurun_out/next_ab_pointwise.log 2>&1; echo "pointwise A/B rc=$?"; grep -E "FAST_ERF|gelu|passed|failed" gpurun_out/next_ab_pointwise.log
timeout 60 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/next_bench.json 2> gpurun_out/next_bench.err; echo "bench rc=$?"; python tools/show_bench.py gpurun_out/next_bench.json | head -3
timeout 400 bash tools/run_ab_overlap.sh > gpurun_out/next_ab_overlap.log 2>&1; echo "overlap A/B rc=$?"; cat gpurun_out/next_ab_overlap.log
