"""Per-kernel timing of csrc/pointwise.cu through the C ABI (CUDA events on the current stream, 10 launches each after
2 warm-ups, buffers rotate so that 240x480 inputs do not sit in L2).  bf16, C = 384.   python tools/bench_pointwise_kernels.py"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from makani_b200 import _ext

lib = _ext.load()
dev = torch.device("cuda:0")
Cn, S = 384, _ext.INSTNORM_SEGMENTS
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
p = lambda t: C.c_void_p(t.data_ptr())                                      # noqa: E731
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for (H, W) in ((240, 480), (721, 1440)):
    hw, nbuf = H * W, (4 if H == 240 else 2)
    xs = [(torch.randn(1, Cn, H, W, device=dev) + 1).to(torch.bfloat16) for _ in range(nbuf)]
    gs = [torch.randn(1, Cn, H, W, device=dev).to(torch.bfloat16) for _ in range(nbuf)]
    out = torch.empty_like(xs[0])
    ws = torch.empty(Cn * S * 2, device=dev)
    stats = torch.empty(Cn, 2, device=dev)
    bsum = torch.zeros(Cn, 2, device=dev)
    gam, bet = torch.ones(Cn, device=dev), torch.zeros(Cn, device=dev)
    nb = Cn * hw * 2
    calls = {
        "stats (1 read)": (1, lambda i: lib.sfno_instnorm_stats(p(xs[i]), _ext.BF16, Cn, hw, 1e-6, p(ws), None, p(stats), st)),
        "apply gelu (1 read 1 write)": (2, lambda i: lib.sfno_instnorm_apply(p(xs[i]), None, _ext.BF16, Cn, Cn, hw, p(stats), p(gam), p(bet), 1, p(out), st)),
        "apply skip (2 reads 1 write)": (3, lambda i: lib.sfno_instnorm_apply(p(xs[i]), p(gs[i]), _ext.BF16, Cn, Cn, hw, p(stats), p(gam), p(bet), 0, p(out), st)),
        "bwd_stats gelu (2 reads)": (2, lambda i: lib.sfno_instnorm_bwd_stats(p(gs[i]), p(xs[i]), _ext.BF16, Cn, Cn, hw, p(stats), p(gam), p(bet), 1, p(ws), st)),
        "bwd_apply gelu (2 reads 1 write)": (3, lambda i: lib.sfno_instnorm_bwd_apply(p(gs[i]), p(xs[i]), _ext.BF16, Cn, Cn, hw, hw, p(stats), p(gam), p(bet), p(bsum), 1, p(out), st)),
        "bias_act gelu (1 read 1 write)": (2, lambda i: lib.sfno_bias_act(p(xs[i]), _ext.BF16, Cn, Cn, hw, p(bet), 1, p(out), st)),
        "bias_act_bwd gelu (2 reads 1 write)": (3, lambda i: lib.sfno_bias_act_bwd(p(gs[i]), p(xs[i]), _ext.BF16, Cn, Cn, hw, p(bet), 1, p(out), p(ws), st)),
        "bias_act gelu in place (1 read 1 write)": (2, lambda i: lib.sfno_bias_act(p(xs[i]), _ext.BF16, Cn, Cn, hw, p(bet), 1, p(xs[i]), st)),
        "bias_act none out of place (1 read 1 write)": (2, lambda i: lib.sfno_bias_act(p(xs[i]), _ext.BF16, Cn, Cn, hw, p(bet), 0, p(out), st)),
        "bias_act none in place (1 read 1 write)": (2, lambda i: lib.sfno_bias_act(p(xs[i]), _ext.BF16, Cn, Cn, hw, p(bet), 0, p(xs[i]), st)),
    }
    for name, (passes, fn) in calls.items():
        for i in range(2):
            assert fn(i % nbuf) == 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(10):
            fn(i % nbuf)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 100
        print(f"{H}x{W} {name:34s} {us:8.1f} us  {passes * nb / us / 1e3:7.1f} GB/s  {passes * nb / us / 1e3 / peak:.2f} of HBM peak")
    del xs, gs
