"""RealSHT / InverseRealSHT: drop-ins for ``torch_harmonics.RealSHT`` / ``InverseRealSHT`` as the reference
uses them (constructed at makani/models/networks/sfnonet.py:536-539, called at
makani/models/common/spectral_convolution.py:131,133,141, makani/utils/losses.py:299-307).

Same constructor, attributes (``nlat nlon lmax mmax grid norm csphase``), non-persistent table buffers
(``weights`` / ``pct``) and forward semantics:

    RealSHT:         x[..., nlat, nlon] real  ->  2*pi*rfft(x, norm="forward")[..., :mmax] contracted with the
                     quadrature-weighted Legendre table  ->  [..., lmax, mmax] complex64
    InverseRealSHT:  the transpose, followed by irfft(n=nlon, norm="forward")

The work is done by hand-written sm_100a kernels (longitude FFT + tcgen05 TF32 Legendre GEMM) through the
C ABI; there is no CPU path.  Gradients are hand-written adjoints built from the same kernels.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from ._nvtx import annotate
from . import ops
from .legendre import legpoly
from .quadrature import clenshaw_curtiss_weights, legendre_gauss_weights, lobatto_weights

__all__ = ["RealSHT", "InverseRealSHT"]


def _grid(nlat: int, grid: str):
    if grid == "legendre-gauss":
        cost, w = legendre_gauss_weights(nlat, -1, 1)
        lmax = nlat
    elif grid == "lobatto":
        cost, w = lobatto_weights(nlat, -1, 1)
        lmax = nlat - 1
    elif grid == "equiangular":
        cost, w = clenshaw_curtiss_weights(nlat, -1, 1)
        lmax = nlat
    else:
        raise ValueError("Unknown quadrature mode")
    return cost, w, lmax


class _TransformBase(nn.Module):
    """Shared plumbing: table precompute (fp64 numpy), lazily created device plan, scale vectors."""

    _table_name = "weights"
    _sp_l_offset = 0          # l_offset handed to the SP <-> complex64 conversions (dense tables: skip nothing)

    def __init__(self, nlat, nlon, lmax=None, mmax=None, grid="lobatto", norm="ortho", csphase=True):
        super().__init__()
        self.nlat, self.nlon, self.grid, self.norm, self.csphase = nlat, nlon, grid, norm, csphase
        cost, w, lmax_default = _grid(nlat, grid)
        self.lmax = lmax or lmax_default
        self.mmax = mmax or nlon // 2 + 1
        self._cost, self._quad_w = cost, w
        self._plans = {}          # device index -> ops.Plan
        self._scales = {}         # (device index, kind) -> tensor

    # subclasses: fp64 numpy table [mmax, lmax, nlat]
    def _table64(self) -> np.ndarray:
        raise NotImplementedError

    def _table_slice(self):
        """(m_offset, table64) actually held by this module (overridden by the azimuth-sharded variants)."""
        return 0, self._host_table64()

    def _host_table64(self) -> np.ndarray:
        """The registered table buffer as fp64 numpy (exact: it is fp64, or the fp32 rounding `.float()` made of it -
        which is what the device plan rounds to anyway)."""
        buf = self._buffers.get(self._table_name)
        if buf is None:
            return self._table64()
        return buf.detach().cpu().double().numpy()

    def _apply(self, fn, *args, **kwargs):
        # The table buffer (`weights` / `pct`: what torch-harmonics registers, non-persistent) follows dtype changes
        # (`.float()`, as makani/models/networks/sfnonet.py:536-539 calls it) but stays in HOST memory when the module
        # moves to a GPU: the kernels read the plan's own device copy (TF32-rounded, padded), and a second copy of the
        # table (333 MB in fp64 at 721x1440) would be dead device memory.  Nothing in the reference reads the buffer.
        name = self._table_name
        tab = self._buffers.pop(name, None)
        try:
            super()._apply(fn, *args, **kwargs)
        finally:
            if tab is not None:
                probe = fn(torch.empty(0, dtype=tab.dtype))
                self._buffers[name] = tab.to(probe.dtype) if probe.dtype.is_floating_point else tab
        return self

    def _plan(self, device: torch.device) -> ops.Plan:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        p = self._plans.get(idx)
        if p is None:
            m_off, tab = self._table_slice()
            p = ops.Plan(self.nlat, self.nlon, self.lmax, tab.shape[0], tab, m_offset=m_off, device=idx)
            self._plans[idx] = p
        return p

    def _scale(self, device: torch.device, kind: str) -> torch.Tensor:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        key = (idx, kind)
        s = self._scales.get(key)
        if s is None:
            mk = {"fwd": ops.fwd_scale, "irfft_adj": ops.irfft_adjoint_scale, "rfft_adj": ops.rfft_adjoint_scale,
                  "one": lambda n, m: torch.ones(m, dtype=torch.float32)}[kind]
            s = mk(self.nlon, self.nlon // 2 + 1).to(device)
            self._scales[key] = s
        return s

    def extra_repr(self):
        return f"nlat={self.nlat}, nlon={self.nlon}, lmax={self.lmax}, mmax={self.mmax}, grid={self.grid}"


def _flatten_spatial(x: torch.Tensor, nlat: int, nlon: int):
    if x.shape[-2] != nlat or x.shape[-1] != nlon:
        raise AssertionError(f"expected [..., {nlat}, {nlon}], got {tuple(x.shape)}")
    lead = x.shape[:-2]
    return x.reshape(1, -1, nlat, nlon), lead


class _SHTFunction(torch.autograd.Function):
    """x [1, BC, nlat, nlon] -> coeffs complex64 [1, BC, lmax, mmax]."""

    @staticmethod
    def forward(ctx, x, mod):
        plan = mod._plan(x.device)
        B, Cn = x.shape[0], x.shape[1]
        sp = ops.sht_to_sp(plan, x, mod._scale(x.device, "fwd"))
        ctx.mod, ctx.in_dtype = mod, x.dtype
        return ops.sp_to_c64(sp, B, Cn, mod.lmax, mod.mmax, l_offset=mod._sp_l_offset)

    @staticmethod
    def backward(ctx, g):
        mod = ctx.mod
        plan = mod._plan(g.device)
        B, Cn = g.shape[0], g.shape[1]
        gsp = ops.c64_to_sp(g.to(torch.complex64), l_offset=mod._sp_l_offset)
        gx = ops.sp_to_field(plan, gsp, B, Cn, mod.nlon, mod._scale(g.device, "rfft_adj"), ctx.in_dtype)
        return gx, None


class _ISHTFunction(torch.autograd.Function):
    """coeffs complex64 [1, BC, lmax, mmax] -> x fp32 [1, BC, nlat, nlon]."""

    @staticmethod
    def forward(ctx, c, mod):
        plan = mod._plan(c.device)
        B, Cn = c.shape[0], c.shape[1]
        sp = ops.c64_to_sp(c, l_offset=mod._sp_l_offset)
        ctx.mod = mod
        return ops.sp_to_field(plan, sp, B, Cn, mod.nlon, mod._scale(c.device, "one"), torch.float32)

    @staticmethod
    def backward(ctx, g):
        mod = ctx.mod
        plan = mod._plan(g.device)
        B, Cn = g.shape[0], g.shape[1]
        gsp = ops.sht_to_sp(plan, g, mod._scale(g.device, "irfft_adj"))
        return ops.sp_to_c64(gsp, B, Cn, mod.lmax, mod.mmax, l_offset=mod._sp_l_offset), None


class RealSHT(_TransformBase):
    """Forward real spherical harmonic transform (analysis)."""

    def __init__(self, nlat, nlon, lmax=None, mmax=None, grid="lobatto", norm="ortho", csphase=True):
        super().__init__(nlat, nlon, lmax, mmax, grid, norm, csphase)
        # same buffer the reference module carries; non-persistent so checkpoints are unaffected
        # (makani/utils/trainer.py:1036-1042)
        self.register_buffer("weights", torch.from_numpy(self._table64()), persistent=False)

    def _table64(self):
        tq = np.flip(np.arccos(self._cost))
        pct = legpoly(self.mmax, self.lmax, tq, norm=self.norm, csphase=self.csphase)
        return np.einsum("mlk,k->mlk", pct, self._quad_w)

    @annotate("RealSHT.forward")
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        ops._need_cuda(x, "RealSHT input")
        x4, lead = _flatten_spatial(x, self.nlat, self.nlon)
        ops.dtype_code(x4.dtype)                         # loud on fp64 etc.
        out = _SHTFunction.apply(x4, self)
        return out.reshape(*lead, self.lmax, self.mmax)


class InverseRealSHT(_TransformBase):
    """Inverse real spherical harmonic transform (synthesis)."""

    _table_name = "pct"

    def __init__(self, nlat, nlon, lmax=None, mmax=None, grid="lobatto", norm="ortho", csphase=True):
        super().__init__(nlat, nlon, lmax, mmax, grid, norm, csphase)
        self.register_buffer("pct", torch.from_numpy(self._table64()), persistent=False)

    def _table64(self):
        t = np.flip(np.arccos(self._cost))
        return legpoly(self.mmax, self.lmax, t, norm=self.norm, inverse=True, csphase=self.csphase)

    @annotate("InverseRealSHT.forward")
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.shape[-2] != self.lmax or x.shape[-1] != self.mmax:
            raise AssertionError(f"expected [..., {self.lmax}, {self.mmax}], got {tuple(x.shape)}")
        if not x.is_complex():
            raise TypeError("InverseRealSHT expects complex coefficients")
        ops._need_cuda(x, "InverseRealSHT input")
        lead = x.shape[:-2]
        c4 = x.reshape(1, -1, self.lmax, self.mmax).to(torch.complex64)
        out = _ISHTFunction.apply(c4, self)
        return out.reshape(*lead, self.nlat, self.nlon)
