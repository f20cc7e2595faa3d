"""Where the whole-model step (tools/benchlib/sfno_model.py) spends its GPU time: torch.profiler (CUPTI) kernel table,
own kernels (sfno::) and library kernels side by side.  usage: python tools/prof_full_model.py [fused|norm|torch]"""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import makani_b200 as mk  # noqa: E402
from benchlib import sfno_model  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "fused"          # fused | norm (fused norms only) | torch
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = sfno_model.SFNO().to(dev)
if mode != "torch":
    mk.fuse_model(model, pointwise=(mode == "fused"))
x = torch.randn(1, 73, 721, 1440, device=dev)


def step():
    for p in model.parameters():
        p.grad = None
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = model(x)
        loss = (y.float() ** 2).mean()
    loss.backward()


for _ in range(2):
    step()
torch.cuda.synchronize()
N = 3
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(N):
        step()
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / N / 1e3, e.count // N) for e in prof.key_averages() if e.device_time_total > 0
        and e.device_type.name == "CUDA"]
rows.sort(key=lambda r: -r[1])
total = sum(r[1] for r in rows)
print(f"# whole-model step, mode = {mode}: sum of kernel time {total:.2f} ms per step")
own = sum(r[1] for r in rows if "sfno" in r[0])
print(f"# own kernels (sfno::) {own:.2f} ms, library / aten kernels {total - own:.2f} ms")
for k, ms, n in rows[:45]:
    print(f"{ms:8.3f} ms  x{n:<4d} {k[:150]}")
