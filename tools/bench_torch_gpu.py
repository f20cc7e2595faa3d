"""The reference-equivalent TORCH path of the spectral step on the GPU ("the real bar to beat" - there is
no reference Blackwell kernel, the reference runs this path as cuFFT + cuBLAS calls from Python): RealSHT =
torch.fft.rfft + two real einsums against the quadrature-weighted Legendre table, dhconv = complex einsum
"bixy,iox->boxy" (contractions.py:131-136), InverseRealSHT = two einsums + torch.fft.irfft, fp32 with TF32 matmuls
allowed (train.py:109-110), input cast `x.float()` and result cast back (spectral_convolution.py:125-146).
Same workload as bench.py: the eight blocks of sfno_linear_73chq_sc3_layers8_edim384_asgl2 chained, forward + backward,
batch 1, bf16 spatial tensors.  Written out here (it does not import oracle/, which is reserved for tests and the CPU
baseline); tables come from the product's fp64 host precompute.

    python tools/bench_torch_gpu.py [--steps 10] [--warmup 3]           # on the GPU box
    python tools/bench_torch_gpu.py --device cpu --small                 # plumbing check without a GPU
Prints one JSON line with ms/step and samples/s."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from makani_b200 import legendre, quadrature


def tables(nlat, nlon, lmax, mmax, grid, dev):
    cost, w = {"legendre-gauss": quadrature.legendre_gauss_weights,
               "equiangular": quadrature.clenshaw_curtiss_weights}[grid](nlat, -1, 1)
    tq = np.flip(np.arccos(cost))
    pct = legendre.legpoly(mmax, lmax, tq)
    weights = torch.from_numpy(np.einsum("mlk,k->mlk", pct, w)).float().to(dev)
    return weights, torch.from_numpy(pct).float().to(dev)


class TorchSHT:
    def __init__(self, nlat, nlon, lmax, mmax, grid, dev):
        self.nlat, self.nlon, self.lmax, self.mmax, self.grid = nlat, nlon, lmax, mmax, grid
        self.weights, self.pct = tables(nlat, nlon, lmax, mmax, grid, dev)

    def sht(self, x):
        x = 2.0 * torch.pi * torch.fft.rfft(x, dim=-1, norm="forward")
        x = torch.view_as_real(x)
        re = torch.einsum("...km,mlk->...lm", x[..., : self.mmax, 0], self.weights)
        im = torch.einsum("...km,mlk->...lm", x[..., : self.mmax, 1], self.weights)
        return torch.complex(re, im)

    def isht(self, c):
        c = torch.view_as_real(c)
        re = torch.einsum("...lm,mlk->...km", c[..., 0], self.pct)
        im = torch.einsum("...lm,mlk->...km", c[..., 1], self.pct)
        return torch.fft.irfft(torch.complex(re, im), n=self.nlon, dim=-1, norm="forward")


def spectral_conv(x, fwd, inv, weight):
    dtype = x.dtype
    residual = x
    x = x.float()
    x = fwd.sht(x).contiguous()
    if fwd is not inv:
        residual = inv.isht(x).to(dtype)
    x = torch.einsum("bixy,iox->boxy", x.contiguous(), weight).contiguous()
    x = inv.isht(x)
    return x.to(dtype), residual


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--device", default="cuda:0")
    ap.add_argument("--small", action="store_true")
    args = ap.parse_args()
    dev = torch.device(args.device)
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
    if args.small:
        C, L, M, big, small = 8, 8, 9, (17, 32, "equiangular"), (8, 16, "legendre-gauss")
    else:
        C, L, M, big, small = 384, 240, 241, (721, 1440, "equiangular"), (240, 480, "legendre-gauss")
    t_big, t_small = TorchSHT(*big[:2], L, M, big[2], dev), TorchSHT(*small[:2], L, M, small[2], dev)
    torch.manual_seed(0)
    ws = []
    for _ in range(8):
        scale = (1.0 / C) ** 0.5 * torch.ones(L, dtype=torch.complex64)
        scale[0] *= 2 ** 0.5
        ws.append((scale * torch.randn(C, C, L, dtype=torch.complex64)).to(dev).requires_grad_(True))
    plan = [(t_big, t_small)] + [(t_small, t_small)] * 6 + [(t_small, t_big)]
    dt = torch.bfloat16
    x0 = torch.rand(1, C, big[0], big[1], device=dev).to(dt).requires_grad_(True)
    g_small = (torch.rand(1, C, small[0], small[1], device=dev) - 0.5).to(dt)
    g_big = (torch.rand(1, C, big[0], big[1], device=dev) - 0.5).to(dt)

    def step():
        x0.grad = None
        for w in ws:
            w.grad = None
        x, outs, gouts = x0, [], []
        for (f, i), w in zip(plan, ws):
            y, res = spectral_conv(x, f, i, w)
            if f is not i:
                outs.append(res)
                gouts.append(g_small if i is t_small else g_big)
            x = y
        outs.append(x)
        gouts.append(g_big)
        torch.autograd.backward(outs, gouts)

    cuda = dev.type == "cuda"
    for _ in range(args.warmup):
        step()
    if cuda:
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    import time
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    if cuda:
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
    else:
        ms = (time.perf_counter() - t0) / args.steps * 1e3
    print(json.dumps({"impl": "torch-gpu restatement (cuFFT + cuBLAS einsums, TF32 allowed)" if cuda else "torch-cpu (plumbing check)",
                      "ms_per_step": ms, "samples_per_s": 1e3 / ms, "steps": args.steps, "warmup": args.warmup,
                      "workload": "8 chained SpectralConv blocks fwd+bwd, batch 1, bf16 spatial tensors",
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9 if cuda else None}))


if __name__ == "__main__":
    main()
