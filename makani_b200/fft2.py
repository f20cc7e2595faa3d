"""RealFFT2 / InverseRealFFT2: drop-ins for the wrappers the reference uses when ``spectral_transform="fft"``
(makani/models/common/layers.py:217-280; selected at makani/models/networks/sfnonet.py:541-553) - SURVEY 8(f) rank 4.

Same constructor (``nlat, nlon, lmax, mmax``), attributes (``lmax mmax truncate lmax_high lmax_low``) and semantics:

    RealFFT2:         y = rfft2(x, norm="ortho"), latitude frequencies [0, ceil(lmax/2)) and the last floor(lmax/2)
                      (the negative ones) concatenated, longitude modes [0, mmax)        -> [..., lmax, mmax] complex64
    InverseRealFFT2:  zero-pads that back to nlat frequencies and applies irfft2(norm="ortho")

Built from the kernels of the spherical transforms: the longitude rFFT is the same bandwidth kernel, and the pruned
latitude DFT (lmax of nlat frequencies) is a dense contraction over the latitudes - exactly the shape of the Legendre
contraction, so it runs on the same tcgen05 TF32 GEMM with the cosine / sine DFT matrices as "tables" that do not depend
on m and have no l >= m structure (``SFNO_M_OFFSET_DENSE`` plans):

    rfft2(x)[f, m]  =  sum_k (cos - i sin)(2 pi f k / nlat) X_m[k] / sqrt(nlat)   =   T_cos(x) - i T_sin(x)
    irfft2(c)       =  S_cos(c) + S_sin(i c)

where T / S are the analysis / synthesis transforms with a real table.  Gradients come from the same hand-written
adjoints as RealSHT / InverseRealSHT.  Like torch's C2R transform, the inverse ignores the imaginary part of the m = 0
(and Nyquist) column after the latitude transform.  Single device; the h/w-sharded variants
(makani/mpu/layers.py:38-169) are not provided.  CUDA only.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn as nn

from . import _ext, ops
from .sht import _ISHTFunction, _SHTFunction, _TransformBase, _flatten_spatial

__all__ = ["RealFFT2", "InverseRealFFT2"]


class _LatitudeDFT(_TransformBase):
    """Analysis (inverse=False) or synthesis (inverse=True) with the table  trig(2 pi f(l) k / nlat) / sqrt(nlat),
    trig = cos or sin, f(l) = the kept latitude frequencies; longitude transform with norm="ortho"."""

    _sp_l_offset = _ext.DENSE_L_OFFSET

    def __init__(self, nlat, nlon, lmax, mmax, trig: str, inverse: bool):
        super().__init__(nlat, nlon, lmax, mmax, grid="equiangular")
        self.trig, self.inverse = trig, inverse
        self._table_name = "pct" if inverse else "weights"

    def _table64(self):
        L, K = self.lmax, self.nlat
        high = math.ceil(L / 2)
        f = np.concatenate([np.arange(high), np.arange(K - (L - high), K)]).astype(np.float64)     # kept frequencies
        ang = 2.0 * np.pi * np.outer(f, np.arange(K)) / K
        t = (np.cos(ang) if self.trig == "cos" else np.sin(ang)) / math.sqrt(K)
        return np.ascontiguousarray(np.broadcast_to(t, (self.mmax, L, K)))

    def _host_table64(self):
        return self._table64()

    def _table_slice(self):
        return _ext.M_OFFSET_DENSE, self._table64()

    def _scale(self, device, kind):
        # the spherical transforms scale the longitude rFFT by 2 pi / nlon ("forward" normalisation); here "ortho"
        idx = device.index if device.index is not None else torch.cuda.current_device()
        key = (idx, kind)
        s = self._scales.get(key)
        if s is None:
            base = super()._scale(device, kind)
            factor = {"fwd": (1.0 / math.sqrt(self.nlon)) / (2.0 * math.pi / self.nlon),
                      "rfft_adj": (1.0 / math.sqrt(self.nlon)) / (2.0 * math.pi / self.nlon),
                      "one": 1.0 / math.sqrt(self.nlon), "irfft_adj": 1.0 / math.sqrt(self.nlon)}[kind]
            s = (base * factor).contiguous()
            self._scales[key] = s
        return s


def _sizes(nlat, nlon, lmax, mmax):
    lmax = min(lmax or nlat, nlat)
    mmax = min(mmax or nlon // 2 + 1, nlon // 2 + 1)
    truncate = not (lmax == nlat and mmax == nlon // 2 + 1)
    return lmax, mmax, truncate


class RealFFT2(nn.Module):
    """layers.py:217-246."""

    def __init__(self, nlat, nlon, lmax=None, mmax=None):
        super().__init__()
        self.nlat, self.nlon = nlat, nlon
        self.lmax, self.mmax, self.truncate = _sizes(nlat, nlon, lmax, mmax)
        self.lmax_high, self.lmax_low = math.ceil(self.lmax / 2), math.floor(self.lmax / 2)
        self._cos = _LatitudeDFT(nlat, nlon, self.lmax, self.mmax, "cos", inverse=False)
        self._sin = _LatitudeDFT(nlat, nlon, self.lmax, self.mmax, "sin", inverse=False)

    def forward(self, x):
        ops._need_cuda(x, "RealFFT2 input")
        x4, lead = _flatten_spatial(x, self.nlat, self.nlon)
        ops.dtype_code(x4.dtype)
        y = _SHTFunction.apply(x4, self._cos) - 1j * _SHTFunction.apply(x4, self._sin)
        return y.reshape(*lead, self.lmax, self.mmax)


class InverseRealFFT2(nn.Module):
    """layers.py:249-280."""

    def __init__(self, nlat, nlon, lmax=None, mmax=None):
        super().__init__()
        self.nlat, self.nlon = nlat, nlon
        self.lmax, self.mmax, self.truncate = _sizes(nlat, nlon, lmax, mmax)
        self.lmax_high, self.lmax_low = math.ceil(self.lmax / 2), math.floor(self.lmax / 2)
        self._cos = _LatitudeDFT(nlat, nlon, self.lmax, self.mmax, "cos", inverse=True)
        self._sin = _LatitudeDFT(nlat, nlon, self.lmax, self.mmax, "sin", inverse=True)

    def forward(self, x):
        if not x.is_complex():
            raise TypeError("InverseRealFFT2 expects complex coefficients")
        ops._need_cuda(x, "InverseRealFFT2 input")
        x = x[..., : self.mmax]                                   # "truncation is implicit but better do it manually"
        if x.shape[-2] != self.lmax:
            raise AssertionError(f"expected [..., {self.lmax}, >= {self.mmax}], got {tuple(x.shape)}")
        lead = x.shape[:-2]
        c4 = x.reshape(1, -1, self.lmax, self.mmax).to(torch.complex64)
        out = _ISHTFunction.apply(c4, self._cos) + _ISHTFunction.apply(1j * c4, self._sin)
        return out.reshape(*lead, self.nlat, self.nlon)
