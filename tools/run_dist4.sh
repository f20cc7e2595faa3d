mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_distributed.py -m gpu -q -x -k "4-1" > gpurun_out/dist4_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/dist4_tests.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29520 tools/prof_spatial.py > gpurun_out/prof_spatial4.log 2>&1; grep "^{" gpurun_out/prof_spatial4.log || tail -20 gpurun_out/prof_spatial4.log
timeout 300 $TR --master-port 29511 tools/bench_spatial.py > gpurun_out/spatial4_p2p.log 2>&1; echo "p2p rc=$?"; grep '^{' gpurun_out/spatial4_p2p.log || tail -20 gpurun_out/spatial4_p2p.log
