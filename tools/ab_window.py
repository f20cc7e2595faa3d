"""A/B of the channel-window mode (MAKANI_B200_WINDOW_MB, SFNO_WINDOW_HINTS are read at import / first use, so one
process per setting):  python tools/ab_window.py  ->  one JSON line: chained step fwd+bwd, forward chain, internal block
forward, per-kernel-family times."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch

from benchlib import workload as W
from benchlib.profiler import KernelProfiler

dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
blocks = W.build_gpu(dev)
dt = torch.bfloat16
grads = W.make_grads(1, dt, dev, chained=True)
x0 = W.make_inputs(1, dt, pinned=False, chained=True)[0].to(dev).requires_grad_(True)


def timed(fn, n, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


out = {"window_mb": os.environ.get("MAKANI_B200_WINDOW_MB", "60"), "hints": os.environ.get("SFNO_WINDOW_HINTS", "1")}
out["step_ms"] = timed(lambda: W.gpu_step_chained(blocks, x0, grads), 10)
plans = []
for blk in blocks:
    plans += [blk.forward_transform._plan(dev), blk.inverse_transform._plan(dev)]
with KernelProfiler(plans) as prof:
    for _ in range(3):
        W.gpu_step_chained(blocks, x0, grads)
    rows = prof.summary()
out["kernels"] = [dict(k=r["kernel"], K=r["shape"].get("K"), n=r["launches"] // 3, ms=round(r["total_ms"] / 3, 3)) for r in rows]
with torch.no_grad():
    out["fwd_chain_ms"] = timed(lambda: W.forward_chain(blocks, x0), 10)
    xs = [torch.rand(1, W.C_EMBED, 240, 480, device=dev).to(dt) for _ in range(4)]
    i = [0]

    def one():
        i[0] += 1
        blocks[1](xs[i[0] % 4])

    out["internal_block_fwd_us"] = timed(one, 24, 4) * 1e3
print(json.dumps(out))
