"""CPU test of the band bookkeeping for nlon = 1440 with more than 241 longitude modes (makani_b200/ops.py:
_dft_fwd_banded / _dft_inv_banded; include/makani_b200.h: sfno_dft_fwd_band, sfno_dft_inv_band, sfno_xm_mirror_conj).

The three C entry points are replaced by numpy models that follow the header's definition of what ONE launch does -
241 bins from three length-480 transforms combined with the band's twiddle table; on the way back bin 0 counted once
(real part), bin 240 doubled (real part) - and read / write the very host buffers the wrappers pass as pointers.  What is
checked is therefore exactly the host logic: which rows each band gets, the mirrored rows 241..479, the per-band scale
slices and end weights, for every truncation that changes the set of launches.  The result must equal numpy's rfft /
irfft of the full 1440-point rows.
"""
import ctypes

import numpy as np
import pytest
import torch

from makani_b200 import _ext, ops

N = 1440
W = np.exp(-2j * np.pi * np.arange(241) / N)
OMEGA = np.exp(-2j * np.pi / 3)
TABLE = {0: W, 1: np.conj(OMEGA) * W, 2: OMEGA * W}


def _floats(ptr, n):
    addr = ptr.value if hasattr(ptr, "value") else int(ptr)
    return np.ctypeslib.as_array((ctypes.c_float * n).from_address(addr))


class FakeLib:
    """numpy stand-ins; XM rows are [2B planes][Cp][Kp] floats, fields [B][Ctot][K][1440] fp32."""

    def __init__(self):
        self.calls = []

    def sfno_dft_fwd_band(self, x, dt, B, Ctot, c0, C, K, band, count, mscale, xm_rows, rnd, st):
        assert dt == _ext.F32 and 1 <= count <= 241
        self.calls.append(("fwd", band, count))
        Cp, Kp = (C + 3) & ~3, (K + 3) & ~3
        xs = _floats(x, B * Ctot * K * N).reshape(B, Ctot, K, N)[:, c0:c0 + C]
        sc = _floats(mscale, count)
        out = _floats(xm_rows, count * 2 * B * Cp * Kp).reshape(count, B, 2, Cp, Kp)
        acc = np.zeros((B, C, K, 241), dtype=complex)
        for r in range(3):
            acc += TABLE[band] ** r * np.fft.rfft(xs[..., r::3].astype(np.float64), axis=-1)
        acc = acc[..., :count] * sc
        out[...] = 0
        out[:, :, 0, :C, :K] = np.moveaxis(acc.real, -1, 0)
        out[:, :, 1, :C, :K] = np.moveaxis(acc.imag, -1, 0)
        return 0

    def sfno_dft_inv_band(self, xm_rows, B, Ctot, c0, C, K, band, count, mscale, y, dt, st):
        assert dt == _ext.F32 and 1 <= count <= 241
        self.calls.append(("inv", band, count))
        Cp, Kp = (C + 3) & ~3, (K + 3) & ~3
        rows = _floats(xm_rows, count * 2 * B * Cp * Kp).reshape(count, B, 2, Cp, Kp)
        sc = _floats(mscale, count)
        H = np.zeros((B, C, K, 241), dtype=complex)
        H[..., :count] = np.moveaxis(rows[:, :, 0, :C, :K] + 1j * rows[:, :, 1, :C, :K], 0, -1) * sc
        ys = _floats(y, B * Ctot * K * N).reshape(B, Ctot, K, N)[:, c0:c0 + C]
        for r in range(3):
            Y = H * np.conj(TABLE[band]) ** r
            Y[..., 0] = Y[..., 0].real
            Y[..., 240] = 2 * Y[..., 240].real
            ys[..., r::3] = np.fft.irfft(Y, 480, axis=-1) * 480
        return 0

    def sfno_xm_mirror_conj(self, src, dst, nrows, nplanes, plane, scale, rnd, st):
        self.calls.append(("mirror", nrows))
        s = _floats(src, nrows * nplanes * plane).reshape(nrows, nplanes, plane)
        d = _floats(dst, nrows * nplanes * plane).reshape(nrows, nplanes, plane)
        sc = _floats(scale, nrows) if scale and getattr(scale, "value", scale) else np.ones(nrows, dtype=np.float32)
        sign = np.where(np.arange(nplanes) % 2 == 1, -1.0, 1.0)[None, :, None]
        d[...] = s[::-1] * sign * sc[:, None, None]
        return 0


@pytest.fixture
def fake(monkeypatch):
    lib = FakeLib()
    monkeypatch.setattr(_ext, "load", lambda: lib)
    monkeypatch.setattr(ops, "_stream", lambda: None)
    return lib


def _decode(xm, B, C, K, M):
    Cp, Kp = (C + 3) & ~3, (K + 3) & ~3
    a = xm.numpy().reshape(M, B, 2, Cp, Kp)
    return np.moveaxis(a[:, :, 0, :C, :K] + 1j * a[:, :, 1, :C, :K], 0, -1)          # [B, C, K, M]


def _encode(c, B, C, K, M):
    Cp, Kp = (C + 3) & ~3, (K + 3) & ~3
    a = np.zeros((M, B, 2, Cp, Kp), dtype=np.float32)
    a[:, :, 0, :C, :K] = np.moveaxis(c.real, -1, 0)
    a[:, :, 1, :C, :K] = np.moveaxis(c.imag, -1, 0)
    return torch.from_numpy(a.reshape(-1))


CASES = [(721, ["fwd0", "fwd1", "mirror239", "fwd2"]), (600, ["fwd0", "fwd1", "mirror239", "fwd2"]),
         (481, ["fwd0", "fwd1", "mirror239", "fwd2"]), (480, ["fwd0", "fwd1", "mirror239"]),
         (400, ["fwd0", "fwd1", "mirror159"]), (242, ["fwd0", "fwd1", "mirror1"])]


@pytest.mark.parametrize("M,launches", CASES)
def test_forward_bands_assemble_the_spectrum(fake, M, launches):
    B, Ctot, c0, C, K = 2, 5, 1, 3, 5                       # a channel window of a larger field, pad channel, Kp = 8
    rng = np.random.default_rng(M)
    x = torch.from_numpy(rng.standard_normal((B, Ctot, K, N)).astype(np.float32))
    mscale = torch.from_numpy((rng.random(721) + 0.5).astype(np.float32))
    xm = torch.zeros(ops.xm_elems(K, M, B, C), dtype=torch.float32)
    ops._dft_fwd_banded(x, Ctot, c0, C, M, mscale, xm, False)
    want = np.fft.rfft(x.numpy()[:, c0:c0 + C].astype(np.float64), axis=-1)[..., :M] * mscale.numpy()[:M]
    np.testing.assert_allclose(_decode(xm, B, C, K, M), want, rtol=0, atol=2e-4 * np.abs(want).max())
    assert [c[0][:3] + str(c[1]) if c[0] != "mirror" else "mirror" + str(c[1]) for c in fake.calls] == launches


@pytest.mark.parametrize("M", [c[0] for c in CASES])
def test_inverse_bands_sum_to_the_inverse_transform(fake, M):
    B, C, K = 1, 2, 3
    rng = np.random.default_rng(1000 + M)
    c = rng.standard_normal((B, C, K, M)) + 1j * rng.standard_normal((B, C, K, M))
    mscale = torch.from_numpy((rng.random(721) + 0.5).astype(np.float32))
    y = ops._dft_inv_banded(_encode(c, B, C, K, M), B, C, K, M, mscale)
    want = np.fft.irfft(c * mscale.numpy()[:M], N, axis=-1) * N      # unnormalised, imaginary parts of DC / Nyquist ignored
    assert y.shape == (B, C, K, N)
    np.testing.assert_allclose(y.numpy(), want, rtol=0, atol=2e-5 * np.abs(want).max())
    assert [k[1] for k in fake.calls if k[0] == "inv"] == ([0, 1, 2] if M > 480 else [0, 1])


def test_banded_only_beyond_241_modes_at_1440(monkeypatch):
    assert ops._banded(1440, 721) and ops._banded(1440, 242)
    assert not ops._banded(1440, 241) and not ops._banded(480, 241) and not ops._banded(2880, 721)
    monkeypatch.setenv("SFNO_FORCE_GENERIC_DFT", "1")
    assert not ops._banded(1440, 721)
