"""Timing of the spectral loss (makani_b200.GeometricH1Loss, relative form: two transforms forward, one backward) on
73 output channels, forward + backward, CUDA events:

  fused      product loss: per-degree energy reduced from the transform's internal tensor (sfno_sp_degree_power)
  formula    the reference's arithmetic (losses.py:327-347) on the complex64 output of the product's RealSHT
  torch      the reference-equivalent path on this GPU: cuFFT rfft + einsum with the fp32 table (TF32 allowed, as
             train.py:109-110 sets it) + the same arithmetic

   python tools/bench_h1_loss.py [iters [nlat [leg]]]   (grids: 240x480 with 240 degrees / 241 modes, 721x1440 with all
   721; nlat / leg restrict the run to one grid / one leg, e.g. `1 721 fused` under ncu)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import makani_b200 as mk

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 5
only_nlat = int(sys.argv[2]) if len(sys.argv) > 2 else None
only_leg = sys.argv[3] if len(sys.argv) > 3 else None
dev = torch.device("cuda:0")
torch.backends.cuda.matmul.allow_tf32 = True
torch.backends.cudnn.allow_tf32 = True


def formula(c):
    p = torch.view_as_real(c)
    p = p[..., 0] ** 2 + p[..., 1] ** 2
    return p[..., :, 0] + 2 * torch.sum(p[..., :, 1:], dim=-1)


def timed(fn):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


for nlat, nlon in ((240, 480), (721, 1440)):
    if only_nlat not in (None, nlat):
        continue
    Cn = 73
    loss = mk.GeometricH1Loss((nlat, nlon)).to(dev)
    sht = loss.sht
    w = loss.h1_weights
    prd = torch.randn(1, Cn, nlat, nlon, device=dev, requires_grad=True)
    tar = torch.randn(1, Cn, nlat, nlon, device=dev)
    table = sht.weights.to(dev, torch.float32)                       # [mmax, lmax, nlat], what torch-harmonics holds

    def norms(e):
        l2, h1 = e.reshape(1, -1).sum(-1), (e * w).reshape(1, -1).sum(-1)
        return 0.5 * torch.sqrt(l2) + 0.5 * torch.sqrt(h1)

    def torch_sht(x):
        xf = torch.view_as_real(2.0 * torch.pi * torch.fft.rfft(x, dim=-1, norm="forward"))[..., : sht.mmax, :]
        return torch.complex(torch.einsum("...km,mlk->...lm", xf[..., 0], table),
                             torch.einsum("...km,mlk->...lm", xf[..., 1], table))

    def run(energy):
        prd.grad = None
        v = (norms(energy(prd - tar)) / norms(energy(tar))).sum()
        v.backward()
        return v

    legs = {"fused": lambda: run(lambda x: mk.losses.degree_power(x, sht)),
            "formula": lambda: run(lambda x: formula(sht(x))),
            "torch": lambda: run(lambda x: formula(torch_sht(x)))}
    vals = {}
    for name, fn in legs.items():
        if only_leg not in (None, name):
            continue
        try:
            vals[name] = float(fn().detach())
            print(f"h1_loss {nlat}x{nlon} C={Cn} lmax={sht.lmax} mmax={sht.mmax} {name:8s} {timed(fn):8.2f} ms  "
                  f"value {vals[name]:.6f}", flush=True)
        except Exception as e:                                        # noqa: BLE001  (report and go on to the next leg)
            print(f"h1_loss {nlat}x{nlon} {name}: FAILED {type(e).__name__}: {str(e)[:200]}", flush=True)
    del loss, sht, table, prd, tar
    torch.cuda.empty_cache()
