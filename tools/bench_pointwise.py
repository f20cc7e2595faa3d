"""Micro-benchmark of the fused instance norm (csrc/pointwise.cu) against the torch modules the reference block
composes (nn.InstanceNorm2d + nn.GELU / + skip add), bf16, C = 384, both grids of the named config.
CUDA events on the current stream, inputs rotate through buffers larger than L2.   python tools/bench_pointwise.py [iters]

GB/s = algorithmic bytes / time with algorithmic bytes = the field read once and written once (forward) resp. g and x
read once, dx written once (backward); the kernels themselves read x twice forward and (g, x) twice backward."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import makani_b200 as mk

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = torch.device("cuda:0")
C = 384
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0


def timeit(fn, nbuf):
    for i in range(3):
        fn(i % nbuf)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i % nbuf)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


for (H, W) in ((240, 480), (721, 1440)):
    nbuf = 4 if H == 240 else 2
    xs = [(torch.randn(1, C, H, W, device=dev) * 2 + 1).to(torch.bfloat16) for _ in range(nbuf)]
    gs = [torch.randn(1, C, H, W, device=dev).to(torch.bfloat16) for _ in range(nbuf)]
    fused = mk.FusedInstanceNorm2d(C, eps=1e-6, affine=True, act="gelu").to(dev)
    ref_n = torch.nn.InstanceNorm2d(C, eps=1e-6, affine=True).to(dev)
    ref_a = torch.nn.GELU()
    nbytes = C * H * W * 2
    with torch.no_grad():
        t_f = timeit(lambda i: fused(xs[i]), nbuf)
        t_r = timeit(lambda i: ref_a(ref_n(xs[i])), nbuf)
    print(f"{H}x{W} norm+gelu fwd : fused {t_f:8.1f} us ({2 * nbytes / t_f / 1e3:7.1f} GB/s, {2 * nbytes / t_f / 1e3 / peak:.2f} of HBM peak)"
          f"   torch {t_r:8.1f} us   x{t_r / t_f:.2f}")

    def fb(mod_fn, i):
        x = xs[i].detach().requires_grad_(True)
        mod_fn(x).backward(gs[i])

    t_f = timeit(lambda i: fb(fused, i), nbuf)
    t_r = timeit(lambda i: fb(lambda x: ref_a(ref_n(x)), i), nbuf)
    print(f"{H}x{W} norm+gelu fwd+bwd: fused {t_f:8.1f} us ({5 * nbytes / t_f / 1e3:7.1f} GB/s, {5 * nbytes / t_f / 1e3 / peak:.2f} of HBM peak)"
          f"   torch {t_r:8.1f} us   x{t_r / t_f:.2f}")
    skip = [torch.randn(1, C, H, W, device=dev).to(torch.bfloat16) for _ in range(nbuf)]
    fused_s = mk.FusedInstanceNorm2d(C, eps=1e-6, affine=True).to(dev)
    with torch.no_grad():
        t_f = timeit(lambda i: fused_s(xs[i], skip[i]), nbuf)
        t_r = timeit(lambda i: ref_n(xs[i]) + skip[i], nbuf)
    print(f"{H}x{W} norm+skip fwd : fused {t_f:8.1f} us ({3 * nbytes / t_f / 1e3:7.1f} GB/s, {3 * nbytes / t_f / 1e3 / peak:.2f} of HBM peak)"
          f"   torch {t_r:8.1f} us   x{t_r / t_f:.2f}")
    del xs, gs, skip
