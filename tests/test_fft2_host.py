"""CPU checks of the host side of makani_b200.fft2 (no kernels run): the cosine / sine latitude-DFT tables and the
frequency selection reproduce the reference wrappers (makani/models/common/layers.py:217-280, restated with numpy's
FFT) when applied to a longitude rFFT in plain numpy, and the rewiring helpers recognise the reference's module shapes."""
import math

import numpy as np
import pytest
import torch
import torch.nn as nn


def _ref_rfft2_truncated(x, nlat, nlon, lmax, mmax):
    y = np.fft.rfft2(x, s=(nlat, nlon), axes=(-2, -1), norm="ortho")
    if lmax == nlat and mmax == nlon // 2 + 1:
        return y
    hi, lo = math.ceil(lmax / 2), math.floor(lmax / 2)
    parts = [y[..., :hi, :mmax]] + ([y[..., nlat - lo:, :mmax]] if lo else [])
    return np.concatenate(parts, axis=-2)


@pytest.mark.parametrize("nlat,nlon,lmax,mmax", [(12, 16, None, None), (13, 20, 8, 7), (16, 24, 9, 10)])
def test_latitude_dft_tables_realise_the_truncated_rfft2(nlat, nlon, lmax, mmax):
    from makani_b200 import fft2
    f = fft2.RealFFT2(nlat, nlon, lmax, mmax)
    L, M = f.lmax, f.mmax
    tc, ts = f._cos._table64(), f._sin._table64()             # [M, L, nlat], identical for every m
    assert tc.shape == (M, L, nlat) and np.array_equal(tc[0], tc[-1])
    rng = np.random.default_rng(0)
    x = rng.standard_normal((3, nlat, nlon))
    X = np.fft.rfft(x, axis=-1, norm="ortho")[..., :M]         # what the longitude kernel hands to the contraction
    y = np.einsum("lk,bkm->blm", tc[0] - 1j * ts[0], X)        # T_cos(x) - i T_sin(x)
    np.testing.assert_allclose(y, _ref_rfft2_truncated(x, nlat, nlon, L, M), atol=1e-12)

    # synthesis tables: S_cos(c) + S_sin(i c) followed by the longitude irfft is irfft2 of the zero-padded spectrum
    g = fft2.InverseRealFFT2(nlat, nlon, lmax, mmax)
    c = rng.standard_normal((3, L, M)) + 1j * rng.standard_normal((3, L, M))
    sc, ss = g._cos._table64()[0], g._sin._table64()[0]
    XM = np.einsum("lk,blm->bkm", sc + 1j * ss, c)
    XM[..., 0] = XM[..., 0].real                               # C2R ignores the imaginary part of the m = 0 column
    if M == nlon // 2 + 1 and nlon % 2 == 0:
        XM[..., -1] = XM[..., -1].real                         # ... and of the Nyquist column
    got = np.fft.irfft(XM, n=nlon, axis=-1, norm="ortho")
    hi, lo = math.ceil(L / 2), math.floor(L / 2)
    pad = np.zeros((3, nlat, nlon // 2 + 1), dtype=complex)
    pad[:, :hi, :M] = c[:, :hi]
    if lo:
        pad[:, nlat - lo:, :M] = c[:, L - lo:]
    want = np.fft.irfft2(pad, s=(nlat, nlon), axes=(-2, -1), norm="ortho")
    np.testing.assert_allclose(got, want, atol=1e-12)


def test_fft2_attributes_follow_the_reference():
    from makani_b200 import fft2
    f = fft2.RealFFT2(721, 1440, 240, 241)
    assert (f.lmax, f.mmax, f.truncate, f.lmax_high, f.lmax_low) == (240, 241, True, 120, 120)
    g = fft2.InverseRealFFT2(32, 64)
    assert (g.lmax, g.mmax, g.truncate) == (32, 33, False)
    with pytest.raises(RuntimeError):                          # no CPU path
        f(torch.zeros(1, 1, 721, 1440))


def test_fuse_pointwise_recognises_the_reference_patterns():
    import makani_b200 as mk
    mlp = nn.Module()
    mlp.fwd = nn.Sequential(nn.Conv2d(4, 8, 1), nn.GELU(), nn.Identity(), nn.Conv2d(8, 4, 1), nn.Identity())
    assert mk.fuse_pointwise(mlp) and mlp._pointwise_convs == (mlp.fwd[0], mlp.fwd[3])
    enc = nn.Module()
    enc.fwd = nn.Sequential(nn.Conv2d(4, 8, 1), nn.GELU(), nn.Conv2d(8, 4, 1, bias=False))
    assert mk.fuse_pointwise(enc)
    assert mk.fuse_pointwise(nn.Conv2d(4, 4, 1, bias=False))
    for other in (nn.Conv2d(4, 4, 3, padding=1), nn.Conv2d(4, 4, 1, stride=2), nn.Linear(4, 4)):
        assert not mk.fuse_pointwise(other)
    drop = nn.Module()
    drop.fwd = nn.Sequential(nn.Conv2d(4, 8, 1), nn.GELU(), nn.Dropout2d(0.1), nn.Conv2d(8, 4, 1))
    assert not mk.fuse_pointwise(drop)
    ckpt = nn.Module()
    ckpt.fwd = nn.Sequential(nn.Conv2d(4, 8, 1), nn.GELU(), nn.Conv2d(8, 4, 1))
    ckpt.checkpointing = 2
    assert not mk.fuse_pointwise(ckpt)
    before = sorted(mlp.state_dict().keys())
    assert before == ["fwd.0.bias", "fwd.0.weight", "fwd.3.bias", "fwd.3.weight"]      # names untouched
