"""Forward-only timing of the spectral path for BASELINE.json config 5 (autoregressive rollout, batch 1, bf16):
the eight SpectralConv blocks chained under `torch.inference_mode()`, packed weights cached (no optimizer between
steps), eager and replayed from a CUDA graph (the reference's inferencer captures its step too).  CUDA events.

    python tools/bench_rollout.py [steps]      -> one JSON line: ms per forward pass of the 8 blocks, eager and graphed"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from makani_b200 import _ext

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
blocks = bench.build_gpu(dev, torch.bfloat16)
x0 = bench.make_inputs(1, torch.bfloat16, pinned=False, chained=True)[0].to(dev)


def forward(x):
    outs = []
    for blk in blocks:
        x, res = blk(x)
        outs.append(res)
    return x, outs[-1]


def timed(fn, n):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


with torch.inference_mode():
    for _ in range(3):
        y, r = forward(x0)
    torch.cuda.synchronize()
    l0 = _ext.launch_count()
    ms_eager = timed(lambda: forward(x0), steps)
    launches = (_ext.launch_count() - l0) // steps

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        forward(x0)
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        yg, rg = forward(x0)
    g.replay()
    torch.cuda.synchronize()
    same = bool(torch.equal(yg, y))
    ms_graph = timed(g.replay, steps)

print(json.dumps({"workload": "8 chained SpectralConv blocks, forward only, batch 1, bf16, cached packed weights",
                  "steps": steps, "ms_eager": ms_eager, "ms_graph": ms_graph, "launches_per_pass": launches,
                  "graph_equals_eager": same}))
