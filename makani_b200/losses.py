"""GeometricH1Loss: the reference's spectral (Sobolev H1) loss on the product's spherical harmonic transform
(makani/utils/losses.py:275-370) - SURVEY 8(f) rank 4, "GeometricH1Loss's SHT".

Same constructor, buffer (``h1_weights`` = l (l + 1)), methods (``abs``, ``rel``, ``forward``) and results as the
reference class.  What differs is how the per-degree energy

    norm2[b, c, l] = |c_l0|^2 + 2 sum_{m >= 1} |c_lm|^2        (losses.py:299-303 and 327-337)

is obtained: the reference materialises the complex64 coefficients [B, C, lmax, mmax] (303 MB for 73 channels at
721 x 1440), squares them and sums over m with four elementwise / reduction passes; here ``degree_power`` reduces the
transform's internal SP tensor directly (``sfno_sp_degree_power``) and its backward scales SP by the incoming gradient
(``sfno_sp_degree_power_bwd``) and runs the transform's adjoint - the coefficient tensor and its gradient never exist.
What is left to torch is arithmetic on [B, C, lmax] values.

At the reference's full resolution the transform keeps all 721 longitude modes; the tiled FFT kernels stop at 241
modes (the model's truncation), so that case runs the general direct-DFT kernel (csrc/dft.cu) - correct, not tuned.
Single device, CUDA only.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn

from . import ops
from ._nvtx import annotate
from .sht import RealSHT, _flatten_spatial


class _DegreePower(torch.autograd.Function):
    """x [1, BC, nlat, nlon] -> fp32 [1, BC, lmax]."""

    @staticmethod
    def forward(ctx, x, mod):
        plan = mod._plan(x.device)
        B, Cn = x.shape[0], x.shape[1]
        sp = ops.sht_to_sp(plan, x, mod._scale(x.device, "fwd"))
        ctx.mod, ctx.in_dtype = mod, x.dtype
        if ctx.needs_input_grad[0]:
            ctx.save_for_backward(sp)
        return ops.sp_degree_power(sp, B, Cn, mod.lmax, mod.mmax)

    @staticmethod
    def backward(ctx, g):
        mod = ctx.mod
        (sp,) = ctx.saved_tensors
        plan = mod._plan(g.device)
        B, Cn = g.shape[0], g.shape[1]
        gsp = ops.sp_degree_power_bwd(sp, g, B, Cn, mod.lmax, mod.mmax)
        gx = ops.sp_to_field(plan, gsp, B, Cn, mod.nlon, mod._scale(g.device, "rfft_adj"), ctx.in_dtype)
        return gx, None


@annotate("degree_power")
def degree_power(x: torch.Tensor, sht: RealSHT) -> torch.Tensor:
    """[..., nlat, nlon] -> fp32 [..., lmax]: energy per degree l of the field's spherical harmonic coefficients, every
    m >= 1 counted twice (its conjugate partner).  Equal to the reference's

        c = torch.view_as_real(sht(x)); c = c[..., 0] ** 2 + c[..., 1] ** 2; c[..., :, 0] + 2 * c[..., :, 1:].sum(-1)
    """
    if not isinstance(sht, RealSHT):
        raise TypeError("makani_b200: degree_power needs a makani_b200.RealSHT")
    if getattr(sht, "comm_size_polar", 1) * getattr(sht, "comm_size_azimuth", 1) > 1:
        raise NotImplementedError("makani_b200: degree_power is single-device (the transform is h/w sharded)")
    ops._need_cuda(x, "degree_power input")
    x4, lead = _flatten_spatial(x, sht.nlat, sht.nlon)
    ops.dtype_code(x4.dtype)
    return _DegreePower.apply(x4, sht).reshape(*lead, sht.lmax)


class GeometricH1Loss(nn.Module):
    """Weighted H1 loss on the sphere; alpha balances the L2 norm and the H1 semi-norm (losses.py:275-370)."""

    def __init__(self, img_shape: Tuple[int, int], p: Optional[float] = 2.0, size_average: Optional[bool] = False,
                 reduction: Optional[bool] = True, absolute: Optional[bool] = False, squared: Optional[bool] = False,
                 alpha: Optional[float] = 0.5):
        super().__init__()
        self.reduction, self.size_average = reduction, size_average
        self.absolute, self.squared, self.alpha = absolute, squared, alpha
        self.sht = RealSHT(*img_shape, grid="equiangular")
        deg = torch.arange(self.sht.lmax, dtype=torch.float32)
        self.register_buffer("h1_weights", deg * (deg + 1))

    def _norms(self, x: torch.Tensor) -> torch.Tensor:
        """alpha |x|_L2 + (1 - alpha) |x|_H1 per example (squared norms if ``squared``)."""
        e = degree_power(x, self.sht)
        n = x.shape[0]
        l2 = e.reshape(n, -1).sum(dim=-1)
        h1 = (e * self.h1_weights).reshape(n, -1).sum(dim=-1)
        if not self.squared:
            l2, h1 = torch.sqrt(l2), torch.sqrt(h1)
        return self.alpha * l2 + (1 - self.alpha) * h1

    def _reduce(self, v: torch.Tensor, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        if not self.reduction:
            return v
        if not self.size_average:
            return torch.sum(v)
        return torch.mean(v) if mask is None else torch.sum(v) / torch.sum(mask)

    def abs(self, prd: torch.Tensor, tar: torch.Tensor) -> torch.Tensor:
        return self._reduce(self._norms(prd - tar))

    def rel(self, prd: torch.Tensor, tar: torch.Tensor, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        v = self._norms(prd - tar) / self._norms(tar)
        if mask is not None:
            v = v * mask
        return self._reduce(v, mask)

    def forward(self, prd: torch.Tensor, tar: torch.Tensor, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        # like the reference, the mask only enters the relative loss
        return self.abs(prd, tar) if self.absolute else self.rel(prd, tar, mask)
