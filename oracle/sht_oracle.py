"""CPU restatement of the reference's SFNO spectral hot path.  TEST INFRASTRUCTURE ONLY.

What is restated, and from where
--------------------------------
* ``SpectralConv.forward``            makani/models/common/spectral_convolution.py:124-148
* ``SpectralConv.__init__`` (init)    makani/models/common/spectral_convolution.py:97-102
* contraction dispatch                makani/models/common/factorizations.py:167-200
* ``_contract_dhconv`` etc.           makani/models/common/contractions.py:121-154
* transform construction / sizes      makani/models/networks/sfnonet.py:503-571
* distributed choreography (twin)     makani/mpu/layers.py:67-102, 134-169
* ``GeometricH1Loss``                 makani/utils/losses.py:275-370
* ``RealSHT`` / ``InverseRealSHT`` / ``DistributedRealSHT`` / quadrature / Legendre tables:
  third-party **torch-harmonics** (pyproject.toml:55 ``torch-harmonics>=0.6.5``;
  docker/Dockerfile:78-79 pins v0.6.4).  The package is NOT vendored under /root/reference and not
  installed here, so these follow its published algorithm (v0.6.5 semantics, the 3-argument
  ``distributed_transpose_*.apply`` form used at makani/mpu/layers.py:73).

Parity pinning status
---------------------
* contractions: PINNED - tests/golden/contractions_*.npz were produced by running the reference's own
  ``contractions.py`` (loaded by file path) in the build container (tools/make_golden.py).
* quadrature: PINNED to the reference's own known-answer test (tests/test_metrics.py:37-45, sum of
  weights) and to numpy's Gauss-Legendre rule.
* spectral loss: PINNED - tests/golden/h1_loss_ref.npz holds values and gradients of the reference's own, unmodified
  ``GeometricH1Loss`` run on this oracle's transform (tools/make_golden.py).
* SHT coefficient / field values: **parity unpinned** - the reference holds no golden vectors for them
  and torch-harmonics cannot be run here.  They are anchored on mathematical invariants instead
  (orthonormality against scipy.special.sph_harm_y, exact round trip, adjointness); see
  tests/test_oracle.py.

Everything here is plain numpy / torch-CPU, fp64 unless a dtype is passed.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch


# --------------------------------------------------------------------------------------------------
# split helper (torch-harmonics distributed/utils.py; modulus distributed/utils.py) - SURVEY 8c
# --------------------------------------------------------------------------------------------------
def compute_split_shapes(size: int, num_chunks: int) -> List[int]:
    if num_chunks == 1:
        return [size]
    chunk = (size + num_chunks - 1) // num_chunks
    last = max(0, size - chunk * (num_chunks - 1))
    if last == 0:
        chunk = size // num_chunks
        last = size - chunk * (num_chunks - 1)
    return [chunk] * (num_chunks - 1) + [last]


# --------------------------------------------------------------------------------------------------
# quadrature (torch-harmonics quadrature.py; used by makani/utils/grids.py:32,77,83)
# --------------------------------------------------------------------------------------------------
def legendre_gauss_weights(n: int, a: float = -1.0, b: float = 1.0) -> Tuple[np.ndarray, np.ndarray]:
    xlg, wlg = np.polynomial.legendre.leggauss(n)
    xlg = (b - a) * 0.5 * xlg + (b + a) * 0.5
    wlg = wlg * (b - a) * 0.5
    return xlg, wlg


def clenshaw_curtiss_weights(n: int, a: float = -1.0, b: float = 1.0) -> Tuple[np.ndarray, np.ndarray]:
    assert n > 1
    tcc = np.cos(np.linspace(np.pi, 0, n))
    if n == 2:
        wcc = np.array([1.0, 1.0])
    else:
        n1 = n - 1
        N = np.arange(1, n1, 2)
        l = len(N)
        m = n1 - l
        v = np.concatenate([2 / N / (N - 2), 1 / N[-1:], np.zeros(m)])
        v = 0 - v[:-1] - v[-1:0:-1]
        g0 = -np.ones(n1)
        g0[l] = g0[l] + n1
        g0[m] = g0[m] + n1
        g = g0 / (n1 ** 2 - 1 + (n1 % 2))
        wcc = np.fft.ifft(v + g).real
        wcc = np.concatenate((wcc, wcc[:1]))
    tcc = (b - a) * 0.5 * tcc + (b + a) * 0.5
    wcc = wcc * (b - a) * 0.5
    return tcc, wcc


def lobatto_weights(n: int, a: float = -1.0, b: float = 1.0, tol: float = 1e-16, maxiter: int = 100):
    tlg = -np.cos(np.pi * np.arange(n) / (n - 1))
    vdm = np.zeros((n, n))
    for _ in range(maxiter):
        tmp = tlg
        vdm[:, 0] = 1.0
        vdm[:, 1] = tlg
        for k in range(2, n):
            vdm[:, k] = ((2 * k - 1) * tlg * vdm[:, k - 1] - (k - 1) * vdm[:, k - 2]) / k
        tlg = tmp - (tlg * vdm[:, n - 1] - vdm[:, n - 2]) / (n * vdm[:, n - 1])
        if np.max(np.abs(tlg - tmp)) < tol:
            break
    wlg = 2.0 / ((n * (n - 1)) * (vdm[:, n - 1] ** 2))
    tlg = (b - a) * 0.5 * tlg + (b + a) * 0.5
    wlg = wlg * (b - a) * 0.5
    return tlg, wlg


def grid_nodes_weights(nlat: int, grid: str):
    """cos(theta) nodes (ascending) + weights, and the default lmax, per torch-harmonics' RealSHT ctor."""
    if grid == "legendre-gauss":
        cost, w = legendre_gauss_weights(nlat, -1, 1)
        lmax_default = nlat
    elif grid == "lobatto":
        cost, w = lobatto_weights(nlat, -1, 1)
        lmax_default = nlat - 1
    elif grid == "equiangular":
        cost, w = clenshaw_curtiss_weights(nlat, -1, 1)
        lmax_default = nlat
    else:
        raise ValueError("Unknown quadrature mode")
    return cost, w, lmax_default


# --------------------------------------------------------------------------------------------------
# Legendre table (torch-harmonics legendre.py: precompute_legpoly) - loop form, fp64
# --------------------------------------------------------------------------------------------------
def precompute_legpoly(mmax: int, lmax: int, t: np.ndarray, norm: str = "ortho", inverse: bool = False,
                       csphase: bool = True) -> np.ndarray:
    """vdm[m, l, k] = orthonormal associated Legendre  P-bar_l^m(cos t_k)  (Condon-Shortley phase)."""
    nmax = max(mmax, lmax)
    vdm = np.zeros((nmax, nmax, len(t)), dtype=np.float64)
    norm_factor = 1.0 if norm == "ortho" else np.sqrt(4 * np.pi)
    norm_factor = 1.0 / norm_factor if inverse else norm_factor
    x = np.cos(t)
    vdm[0, 0, :] = norm_factor / np.sqrt(4 * np.pi)
    for l in range(1, nmax):
        vdm[l - 1, l, :] = np.sqrt(2 * l + 1) * x * vdm[l - 1, l - 1, :]
        vdm[l, l, :] = np.sqrt((2 * l + 1) * (1 + x) * (1 - x) / 2 / l) * vdm[l - 1, l - 1, :]
    for l in range(2, nmax):
        for m in range(0, l - 1):
            vdm[m, l, :] = (x * np.sqrt((2 * l - 1) / (l - m) * (2 * l + 1) / (l + m)) * vdm[m, l - 1, :]
                            - np.sqrt((l + m - 1) / (l - m) * (2 * l + 1) / (2 * l - 3) * (l - m - 1) / (l + m))
                            * vdm[m, l - 2, :])
    if norm == "schmidt":
        for l in range(0, nmax):
            if inverse:
                vdm[:, l, :] = vdm[:, l, :] * np.sqrt(2 * l + 1)
            else:
                vdm[:, l, :] = vdm[:, l, :] / np.sqrt(2 * l + 1)
    vdm = vdm[:mmax, :lmax]
    if csphase:
        for m in range(1, mmax, 2):
            vdm[m] *= -1
    return vdm


# --------------------------------------------------------------------------------------------------
# transforms (torch-harmonics sht.py RealSHT / InverseRealSHT)
# --------------------------------------------------------------------------------------------------
class RealSHT:
    """x[..., nlat, nlon] real -> c[..., lmax, mmax] complex.  Call sites: sfnonet.py:536,538."""

    def __init__(self, nlat, nlon, lmax=None, mmax=None, grid="lobatto", norm="ortho", csphase=True,
                 dtype=torch.float64):
        self.nlat, self.nlon, self.grid, self.norm, self.csphase = nlat, nlon, grid, norm, csphase
        cost, w, lmax_default = grid_nodes_weights(nlat, grid)
        self.lmax = lmax or lmax_default
        self.mmax = mmax or nlon // 2 + 1
        tq = np.flip(np.arccos(cost))
        pct = precompute_legpoly(self.mmax, self.lmax, tq, norm=norm, csphase=csphase)
        weights = np.einsum("mlk,k->mlk", pct, w)
        self.weights_f64 = weights
        self.weights = torch.from_numpy(weights).to(dtype)

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        assert x.shape[-2] == self.nlat and x.shape[-1] == self.nlon
        x = 2.0 * torch.pi * torch.fft.rfft(x, dim=-1, norm="forward")
        x = torch.view_as_real(x)
        wt = self.weights.to(x.dtype)
        re = torch.einsum("...km,mlk->...lm", x[..., : self.mmax, 0], wt)
        im = torch.einsum("...km,mlk->...lm", x[..., : self.mmax, 1], wt)
        return torch.complex(re, im)


class InverseRealSHT:
    """c[..., lmax, mmax] complex -> x[..., nlat, nlon] real.  Call sites: sfnonet.py:537,539."""

    def __init__(self, nlat, nlon, lmax=None, mmax=None, grid="lobatto", norm="ortho", csphase=True,
                 dtype=torch.float64):
        self.nlat, self.nlon, self.grid, self.norm, self.csphase = nlat, nlon, grid, norm, csphase
        cost, _, lmax_default = grid_nodes_weights(nlat, grid)
        self.lmax = lmax or lmax_default
        self.mmax = mmax or nlon // 2 + 1
        t = np.flip(np.arccos(cost))
        pct = precompute_legpoly(self.mmax, self.lmax, t, norm=norm, inverse=True, csphase=csphase)
        self.pct_f64 = pct
        self.pct = torch.from_numpy(pct).to(dtype)

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        assert x.shape[-2] == self.lmax and x.shape[-1] == self.mmax
        x = torch.view_as_real(x)
        pct = self.pct.to(x.dtype)
        rl = torch.einsum("...lm,mlk->...km", x[..., 0], pct)
        im = torch.einsum("...lm,mlk->...km", x[..., 1], pct)
        xs = torch.complex(rl, im)
        return torch.fft.irfft(xs, n=self.nlon, dim=-1, norm="forward")


# --------------------------------------------------------------------------------------------------
# channel contractions (makani/models/common/contractions.py:121-154)
# --------------------------------------------------------------------------------------------------
def contract_diagonal(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:      # contractions.py:122-127
    return torch.einsum("bixy,ioxy->boxy", a, b)


def contract_dhconv(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:        # contractions.py:131-136
    return torch.einsum("bixy,iox->boxy", a, b)


def contract_sep_diagonal(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:  # contractions.py:140-145
    return torch.einsum("bixy,ixy->boxy", a, b)


def contract_sep_dhconv(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:    # contractions.py:149-154
    return torch.einsum("bixy,ix->boxy", a, b)


def contract(x, weight, separable=False, operator_type="diagonal"):
    """factorizations.py:167-200 (_contract_dense_pytorch, complex=True branch)."""
    x = x.contiguous()
    if separable:
        if operator_type == "diagonal":
            x = contract_sep_diagonal(x, weight)
        elif operator_type == "dhconv":
            x = contract_sep_dhconv(x, weight)
        else:
            raise ValueError(f"Unkonw operator type {operator_type}")
    else:
        if operator_type == "diagonal":
            x = contract_diagonal(x, weight)
        elif operator_type == "dhconv":
            x = contract_dhconv(x, weight)
        else:
            raise ValueError(f"Unkonw operator type {operator_type}")
    return x.contiguous()


def init_spectral_weight(in_channels, out_channels, modes_lat, modes_lon=None, operator_type="dhconv",
                         separable=False, gain=1.0, generator: Optional[torch.Generator] = None):
    """spectral_convolution.py:66-102: weight shape and init (scale*randn, l=0 scaled by sqrt(2))."""
    shape = [in_channels] + ([] if separable else [out_channels])
    if operator_type == "diagonal":
        shape += [modes_lat, modes_lon]
    elif operator_type == "dhconv":
        shape += [modes_lat]
    else:
        raise ValueError(f"Unsupported operator type f{operator_type}")
    # the reference's length-L scale broadcasts over the last axis: l for dhconv, m for diagonal (needs L == M there)
    scale = math.sqrt(gain / in_channels) * torch.ones(modes_lat, dtype=torch.complex64)
    scale[0] *= math.sqrt(2.0)
    return scale * torch.randn(*shape, dtype=torch.complex64, generator=generator)


def spectral_conv_forward(x, forward_transform, inverse_transform, weight, operator_type="dhconv",
                          separable=False, bias=None, compute_dtype=torch.float32):
    """spectral_convolution.py:124-148.  Returns (y, residual)."""
    dtype = x.dtype
    residual = x
    x = x.to(compute_dtype)
    scale_residual = (forward_transform.nlat != inverse_transform.nlat
                      or forward_transform.nlon != inverse_transform.nlon
                      or forward_transform.grid != inverse_transform.grid)
    x = forward_transform(x).contiguous()
    if scale_residual:
        residual = inverse_transform(x).to(dtype)
    cdt = torch.complex128 if compute_dtype == torch.float64 else torch.complex64
    x = contract(x, weight.to(cdt), separable=separable, operator_type=operator_type)
    x = inverse_transform(x)
    if bias is not None:
        x = x + bias
    return x.to(dtype), residual


# --------------------------------------------------------------------------------------------------
# distributed SHT: what each (h, w) rank must hold.  The reference choreography (4 all-to-alls,
# SURVEY 3.3 / mpu/layers.py:67-102) is a pure re-sharding, so the per-rank result is a slice of the
# single-process result.
# --------------------------------------------------------------------------------------------------
def shard(x: torch.Tensor, dim: int, sizes: Sequence[int], rank: int) -> torch.Tensor:
    start = sum(sizes[:rank])
    return x.narrow(dim, start, sizes[rank])


def distributed_sht_expected(x_full, sht: RealSHT, h: int, w: int, rank_h: int, rank_w: int):
    c = sht(x_full)
    l_shapes = compute_split_shapes(sht.lmax, h)
    m_shapes = compute_split_shapes(sht.mmax, w)
    return shard(shard(c, -2, l_shapes, rank_h), -1, m_shapes, rank_w)


def distributed_isht_expected(c_full, isht: InverseRealSHT, h: int, w: int, rank_h: int, rank_w: int):
    x = isht(c_full)
    lat_shapes = compute_split_shapes(isht.nlat, h)
    lon_shapes = compute_split_shapes(isht.nlon, w)
    return shard(shard(x, -2, lat_shapes, rank_h), -1, lon_shapes, rank_w)


# --------------------------------------------------------------------------------------------------
# spectral loss (makani/utils/losses.py:275-370)
# --------------------------------------------------------------------------------------------------
def degree_power(x: torch.Tensor, sht: RealSHT) -> torch.Tensor:
    """losses.py:299-301 / 327-329: |c|^2 per (l, m), column m = 0 once and every other column twice -> [..., lmax]."""
    c = sht(x)
    p = c.real ** 2 + c.imag ** 2
    return p[..., :, 0] + 2 * torch.sum(p[..., :, 1:], dim=-1)


def geometric_h1_loss(prd, tar, sht: RealSHT, mask=None, alpha=0.5, absolute=False, squared=False, reduction=True,
                      size_average=False):
    """GeometricH1Loss.forward (losses.py:365-370) with its abs (:296-322) and rel (:324-363) branches."""
    n = prd.shape[0]
    h1w = torch.arange(sht.lmax, dtype=prd.dtype)
    h1w = h1w * (h1w + 1)                                                     # losses.py:293-294

    def norms(x):
        e = degree_power(x, sht)
        l2 = e.reshape(n, -1).sum(dim=-1)                                     # :302 / :330
        h1 = (e * h1w).reshape(n, -1).sum(dim=-1)                             # :303 / :331
        if not squared:
            return alpha * torch.sqrt(l2) + (1 - alpha) * torch.sqrt(h1)      # :306 / :340-341
        return alpha * l2 + (1 - alpha) * h1                                  # :308 / :343-344

    if absolute:
        val = norms(prd - tar)
        mask = None                                                           # abs() takes no mask (:365-367)
    else:
        val = norms(prd - tar) / norms(tar)                                   # :347
        if mask is not None:
            val = val * mask
    if not reduction:
        return val
    if size_average:
        return torch.mean(val) if mask is None else torch.sum(val) / torch.sum(mask)
    return torch.sum(val)


def rel_l2(a: torch.Tensor, b: torch.Tensor) -> float:
    """tests/distributed/tests_fft.py:231 metric: mean over leading dims of ||a-b||/||b|| on last 2 dims."""
    a = a.to(torch.complex128) if a.is_complex() else a.to(torch.float64)
    b = b.to(a.dtype)
    num = torch.linalg.vector_norm(a - b, dim=(-2, -1))
    den = torch.linalg.vector_norm(b, dim=(-2, -1)).clamp_min(1e-300)
    return float((num / den).mean())
