"""The reference-equivalent TORCH path of the spectral step on the GPU ("the real bar to beat" - there is
no reference Blackwell kernel, the reference runs this path as cuFFT + cuBLAS calls from Python): RealSHT =
torch.fft.rfft + two real einsums against the quadrature-weighted Legendre table, dhconv = complex einsum
"bixy,iox->boxy" (contractions.py:131-136), InverseRealSHT = two einsums + torch.fft.irfft, fp32 with TF32 matmuls
allowed (train.py:109-110), input cast `x.float()` and result cast back (spectral_convolution.py:125-146).
Same workload as bench.py: the eight blocks of sfno_linear_73chq_sc3_layers8_edim384_asgl2 chained, forward + backward,
batch 1, bf16 spatial tensors.  Written out here (it does not import oracle/, which is reserved for tests and the CPU
baseline); tables come from the product's fp64 host precompute.

bench.py runs `measure()` in the same process and prints the result as `gpu_baseline`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
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



def measure(dev, steps=5, warmup=2, small=False, dtype=torch.bfloat16):
    """ms per chained 8-block step (fwd + bwd) of the torch restatement on `dev`; frees its tensors afterwards."""
    import time
    tf32 = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = True          # makani/train.py:109-110
    torch.backends.cudnn.allow_tf32 = True
    if small:
        C, L, M, big, small_g = 8, 8, 9, (17, 32, "equiangular"), (8, 16, "legendre-gauss")
    else:
        C, L, M, big, small_g = 384, 240, 241, (721, 1440, "equiangular"), (240, 480, "legendre-gauss")
    t_big, t_small = TorchSHT(*big[:2], L, M, big[2], dev), TorchSHT(*small_g[:2], L, M, small_g[2], dev)
    torch.manual_seed(0)
    ws = []
    for _ in range(8):
        scale = (1.0 / C) ** 0.5 * torch.ones(L, dtype=torch.complex64)
        scale[0] *= 2 ** 0.5
        ws.append((scale * torch.randn(C, C, L, dtype=torch.complex64)).to(dev).requires_grad_(True))
    plan = [(t_big, t_small)] + [(t_small, t_small)] * 6 + [(t_small, t_big)]
    x0 = torch.rand(1, C, big[0], big[1], device=dev).to(dtype).requires_grad_(True)
    g_small = (torch.rand(1, C, small_g[0], small_g[1], device=dev) - 0.5).to(dtype)
    g_big = (torch.rand(1, C, big[0], big[1], device=dev) - 0.5).to(dtype)

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

    cuda = torch.device(dev).type == "cuda"
    if cuda:
        torch.cuda.reset_peak_memory_stats()
    for _ in range(warmup):
        step()
    if cuda:
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    if cuda:
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        peak = torch.cuda.max_memory_allocated() / 1e9
    else:
        ms = (time.perf_counter() - t0) / steps * 1e3
        peak = None
    torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = tf32
    out = {"impl": "torch restatement of the reference path on this GPU: cuFFT rfft/irfft + cuBLAS einsums, fp32 with TF32 "
                   "allowed (train.py:109-110), x.float() / .to(dtype) casts (spectral_convolution.py:124-148)",
           "ms_per_step": ms, "value": 1e3 / ms, "unit": "samples/s", "steps": steps, "warmup": warmup,
           "workload": "same chained 8-block step, fwd+bwd, batch 1, bf16 spatial tensors", "peak_mem_gb": peak}
    del ws, x0, g_small, g_big, t_big, t_small, plan
    if cuda:
        torch.cuda.empty_cache()
    return out
