"""Micro-benchmark of the longitude DFT kernels (CUDA events, inputs > L2 rotate through several buffers).
   python tools/bench_dft.py [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from makani_b200 import ops

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = torch.device("cuda:0")
C, M = 384, 241
for (K, W) in ((240, 480), (721, 1440)):
    nbuf = 4 if K == 240 else 2
    xs = [torch.rand(1, C, K, W, device=dev).to(torch.bfloat16) for _ in range(nbuf)]
    sc = ops.fwd_scale(W, W // 2 + 1).to(dev)
    one = torch.ones(W // 2 + 1, device=dev)
    xms = [ops.dft_fwd(x, M, sc) for x in xs]
    for name, fn in (("fwd", lambda i: ops.dft_fwd(xs[i % nbuf], M, sc)),
                     ("inv", lambda i: ops.dft_inv(xms[i % nbuf], 1, C, K, W, M, one, torch.bfloat16))):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / iters * 1e3
        gb = C * K * (W * 2 + M * 8) / 1e9
        print(f"dft_{name} {K}x{W}: {us:8.1f} us  {gb / us * 1e6:7.1f} GB/s  env={os.environ.get('SFNO_P2_OCC','')},{os.environ.get('SFNO_DFT_IMPL','')}")
