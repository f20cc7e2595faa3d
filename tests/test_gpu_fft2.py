"""GPU parity tests of makani_b200.RealFFT2 / InverseRealFFT2 against the reference wrappers
(makani/models/common/layers.py:217-280), restated here with torch.fft in fp64: rfft2(norm="ortho") + the truncation to
the lowest lmax latitude frequencies (positive and negative) and mmax longitude modes, and its inverse.

Tolerance: TOL_T = 1e-3 relative L2, as for the spherical transforms (the latitude DFT is one TF32 contraction with
TF32-rounded cosine / sine matrices); gradients are compared with autograd through the fp64 restatement.
"""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL_T = 1e-3

CASES = [  # nlat, nlon, lmax, mmax
    (32, 64, None, None),        # no truncation (Nyquist column included)
    (33, 64, 20, 17),            # odd nlat, truncated both ways, even lmax
    (24, 48, 13, 12),            # odd lmax: ceil / floor split of the kept frequencies
    (60, 480, 40, 41),           # nlon = 480: the fast longitude kernels
]


@pytest.fixture(scope="module")
def mk(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import makani_b200
    return makani_b200


def _rel(a, b):
    a, b = (torch.view_as_real(a), torch.view_as_real(b)) if a.is_complex() else (a, b)
    return float((a.double() - b.double()).norm() / b.double().norm())


def ref_fwd(x, nlat, nlon, lmax, mmax):
    """layers.py:239-246"""
    y = torch.fft.rfft2(x, s=(nlat, nlon), dim=(-2, -1), norm="ortho")
    if not (lmax == nlat and mmax == nlon // 2 + 1):
        hi, lo = math.ceil(lmax / 2), math.floor(lmax / 2)
        y = torch.cat((y[..., :hi, :mmax], y[..., nlat - lo:, :mmax]), dim=-2) if lo else y[..., :hi, :mmax]
    return y


def ref_inv(c, nlat, nlon, lmax, mmax):
    """layers.py:268-280"""
    xt = c[..., :mmax]
    if not (lmax == nlat and mmax == nlon // 2 + 1):
        hi, lo = math.ceil(lmax / 2), math.floor(lmax / 2)
        xth = F.pad(xt[..., :hi, :], (0, 0, 0, nlat - lmax))
        xt = torch.cat([xth, xt[..., lmax - lo:, :]], dim=-2) if lo else xth
    return torch.fft.irfft2(xt, s=(nlat, nlon), dim=(-2, -1), norm="ortho")


@pytest.mark.parametrize("nlat,nlon,lmax,mmax", CASES)
def test_real_fft2_forward_backward(mk, nlat, nlon, lmax, mmax):
    torch.manual_seed(0)
    mod = mk.RealFFT2(nlat, nlon, lmax, mmax).to(DEV)
    L, M = mod.lmax, mod.mmax
    x = torch.randn(2, 3, nlat, nlon, device=DEV, requires_grad=True)
    y = mod(x)
    assert y.shape == (2, 3, L, M) and y.dtype == torch.complex64
    xr = x.detach().double().requires_grad_(True)
    yr = ref_fwd(xr, nlat, nlon, L, M)
    assert _rel(y.detach(), yr.detach()) < TOL_T
    g = torch.randn(2, 3, L, M, dtype=torch.complex64, device=DEV)
    (y * g.conj()).real.sum().backward()
    (yr * g.to(torch.complex128).conj()).real.sum().backward()
    assert _rel(x.grad, xr.grad) < TOL_T


@pytest.mark.parametrize("nlat,nlon,lmax,mmax", CASES)
def test_inverse_real_fft2_forward_backward(mk, nlat, nlon, lmax, mmax):
    torch.manual_seed(1)
    mod = mk.InverseRealFFT2(nlat, nlon, lmax, mmax).to(DEV)
    L, M = mod.lmax, mod.mmax
    c = torch.randn(2, 3, L, M, dtype=torch.complex64, device=DEV, requires_grad=True)
    x = mod(c)
    assert x.shape == (2, 3, nlat, nlon) and x.dtype == torch.float32
    cr = c.detach().to(torch.complex128).requires_grad_(True)
    xr = ref_inv(cr, nlat, nlon, L, M)
    assert _rel(x.detach(), xr.detach()) < TOL_T
    g = torch.randn(2, 3, nlat, nlon, device=DEV)
    (x * g).sum().backward()
    (xr * g.double()).sum().backward()
    assert _rel(c.grad, cr.grad) < TOL_T


def test_fft2_round_trip_and_bf16_input(mk):
    """Untruncated: inverse(forward(x)) = x; a bf16 field is accepted like the spherical transforms accept it."""
    nlat, nlon = 32, 64
    f, i = mk.RealFFT2(nlat, nlon).to(DEV), mk.InverseRealFFT2(nlat, nlon).to(DEV)
    x = torch.randn(1, 4, nlat, nlon, device=DEV)
    assert _rel(i(f(x)), x) < 2 * TOL_T
    xb = x.to(torch.bfloat16)
    assert _rel(f(xb), ref_fwd(xb.double(), nlat, nlon, nlat, nlon // 2 + 1)) < TOL_T
    with pytest.raises(RuntimeError):
        f(x.cpu())


@pytest.mark.parametrize("op", ["dhconv", "diagonal"])
def test_spectral_conv_on_fft2_transforms(mk, op):
    """spectral_transform="fft" (sfnonet.py:541-553): SpectralConv between RealFFT2 and InverseRealFFT2.  Fourier
    coefficients have no m <= l structure, so the contraction must be dense; compared with the reference's forward
    (spectral_convolution.py:124-148) restated with torch.fft and a complex einsum in fp64.  2e-3 (TOL_BLOCK)."""
    torch.manual_seed(2)
    nlat, nlon, L, M, C = 30, 60, 20, 21, 6
    fwd, inv = mk.RealFFT2(nlat, nlon, L, M).to(DEV), mk.InverseRealFFT2(nlat, nlon, L, M).to(DEV)
    conv = mk.SpectralConv(fwd, inv, C, C, operator_type=op, bias=False, gain=1.0).to(DEV)
    x = torch.randn(2, C, nlat, nlon, device=DEV, requires_grad=True)
    g = torch.randn(2, C, nlat, nlon, device=DEV)
    y, res = conv(x)
    assert res is x or torch.equal(res, x)
    y.backward(g)
    xr = x.detach().double().requires_grad_(True)
    wr = conv.weight.detach().to(torch.complex128).requires_grad_(True)
    eq = "bixy,iox->boxy" if op == "dhconv" else "bixy,ioxy->boxy"
    yr = ref_inv(torch.einsum(eq, ref_fwd(xr, nlat, nlon, L, M), wr), nlat, nlon, L, M)
    yr.backward(g.double())
    assert _rel(y.detach(), yr.detach()) < 2e-3
    assert _rel(x.grad, xr.grad) < 2e-3
    assert _rel(conv.weight.grad, wr.grad) < 2e-3
