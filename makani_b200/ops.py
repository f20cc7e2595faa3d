"""Thin torch-tensor wrappers over the C ABI (one function per kernel family).

Device memory, streams and dtypes come from torch; the arithmetic is all in libmakani_b200.so.
Every wrapper allocates its output with ``torch.empty`` on the input's device and enqueues on
``torch.cuda.current_stream()`` - no host synchronisation, so calls are CUDA-graph capturable
after one warm-up call per shape (tile schedules / twiddle tables are built on first use).

Internal tensors (see include/makani_b200.h):
    XM  flat fp32  [mmax][RS][Kp]      SP  flat fp32  [lmax][mmax][RS]      WP  flat fp32 [lmax][2][Coutp][Cinp]
"""
from __future__ import annotations

import ctypes as C
import math
import os

import numpy as np
import torch

from . import _ext

_DTYPES = {torch.float32: _ext.F32, torch.bfloat16: _ext.BF16, torch.float16: _ext.F16}


def _rup(x, m):
    return (x + m - 1) // m * m


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t: torch.Tensor):
    return C.c_void_p(t.data_ptr())


def _need_cuda(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise RuntimeError(f"makani_b200: {name} must be a CUDA tensor (got {t.device}); this package has no CPU path")


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DTYPES[dt]
    except KeyError:
        raise TypeError(f"makani_b200: unsupported dtype {dt}; supported: float32, bfloat16, float16") from None


class Plan:
    """Owns one device Legendre table [mmax][lmax][nlat] (``sfno_plan_create``)."""

    def __init__(self, nlat, nlon, lmax, mmax, table64: np.ndarray, m_offset: int = 0, device=None):
        lib = _ext.load()
        table64 = np.ascontiguousarray(table64, dtype=np.float64)
        assert table64.shape == (mmax, lmax, nlat), (table64.shape, (mmax, lmax, nlat))
        self.nlat, self.nlon, self.lmax, self.mmax, self.m_offset = nlat, nlon, lmax, mmax, m_offset
        self.device = torch.device(device if device is not None else torch.cuda.current_device())
        self._h = C.c_void_p()
        with torch.cuda.device(self.device):
            _ext.check(lib.sfno_plan_create(nlat, nlon, lmax, mmax, m_offset,
                                            table64.ctypes.data_as(C.c_void_p), C.byref(self._h)), "sfno_plan_create")

    @property
    def handle(self):
        return self._h

    def __del__(self):
        try:                                   # module globals may already be gone at interpreter shutdown
            h = getattr(self, "_h", None)
            if h is not None and h.value:
                _ext.load().sfno_plan_destroy(h)
                self._h = None
        except Exception:
            pass


def xm_elems(nlat, mmax, B, Cn):
    return mmax * 2 * B * _rup(Cn, 4) * _rup(nlat, 4)


def sp_elems(lmax, mmax, B, Cn):
    return lmax * mmax * 2 * B * _rup(Cn, 4)


def wp_elems(lmax, Cin, Cout):
    return lmax * 2 * _rup(Cout, 4) * _rup(Cin, 4)


# ------------------------------------------------------------------------------------------ DFT
def fwd_scale(nlon: int, mmax: int) -> torch.Tensor:
    """RealSHT: X = 2*pi * rfft(x, norm="forward")  ->  2*pi/nlon on every bin."""
    return torch.full((mmax,), 2.0 * math.pi / nlon, dtype=torch.float32)


def irfft_adjoint_scale(nlon: int, mmax: int) -> torch.Tensor:
    """Adjoint of irfft(n=nlon, norm="forward"): bins are doubled except m = 0 and Nyquist."""
    s = torch.full((mmax,), 2.0, dtype=torch.float32)
    s[0] = 1.0
    if nlon % 2 == 0 and mmax > nlon // 2:
        s[nlon // 2] = 1.0
    return s


def rfft_adjoint_scale(nlon: int, mmax: int) -> torch.Tensor:
    """Adjoint of 2*pi*rfft(norm="forward") expressed through the irfft kernel (which doubles bins 0<m<Nyq)."""
    s = torch.full((mmax,), 0.5, dtype=torch.float32)
    s[0] = 1.0
    if nlon % 2 == 0 and mmax > nlon // 2:
        s[nlon // 2] = 1.0
    return s * (2.0 * math.pi / nlon)


# nlon = 1440 with more than 241 modes (the spectral loss's transform keeps all 721): the tiled kernels produce 241 bins
# per launch, three launches with modified twiddle tables cover the spectrum (include/makani_b200.h: sfno_dft_fwd_band).
def _banded(nlon: int, mmax: int) -> bool:
    return nlon == 1440 and mmax > 241 and os.environ.get("SFNO_FORCE_GENERIC_DFT", "0") != "1"


def _fptr(t: torch.Tensor, offset_floats: int = 0):
    return C.c_void_p(t.data_ptr() + 4 * offset_floats)


def _dft_fwd_banded(x: torch.Tensor, Ctot: int, c0: int, c: int, M: int, mscale: torch.Tensor, xm: torch.Tensor,
                    round_tf32: bool):
    """Channels [c0, c0 + c) of x [B, Ctot, nlat, 1440] -> compact XM (M rows).  Bins 0..240 and 480.. come straight
    from the kernel; 241..479 are the mirrored conjugates of a third launch that writes a scratch."""
    lib, st, dt = _ext.load(), _stream(), dtype_code(x.dtype)
    B, K = x.shape[0], x.shape[2]
    plane = _rup(c, 4) * _rup(K, 4)
    row = 2 * B * plane

    def band(b, count, scale, scale_off, dst, dst_off, rnd):
        _ext.check(lib.sfno_dft_fwd_band(_ptr(x), dt, B, Ctot, c0, c, K, b, count, _fptr(scale, scale_off),
                                         _fptr(dst, dst_off), int(rnd), st), "sfno_dft_fwd_band")

    band(0, 241, mscale, 0, xm, 0, round_tf32)
    n1 = min(M, 480) - 241                                   # bins 241 .. 240 + n1 = 480 - k, k = 239 .. 240 - n1
    if n1 > 0:
        tmp = torch.empty(240 * row, dtype=torch.float32, device=x.device)
        ones = torch.ones(240, dtype=torch.float32, device=x.device)
        band(1, 240, ones, 0, tmp, 0, False)
        _ext.check(lib.sfno_xm_mirror_conj(_fptr(tmp, (240 - n1) * row), _fptr(xm, 241 * row), n1, 2 * B, plane,
                                           _fptr(mscale, 241), int(round_tf32), st), "sfno_xm_mirror_conj")
    if M > 480:
        band(2, M - 480, mscale, 480, xm, 480 * row, round_tf32)


def _dft_inv_banded(xm: torch.Tensor, B: int, c: int, K: int, M: int, mscale: torch.Tensor) -> torch.Tensor:
    """Compact XM (M rows, c channels) -> fp32 [B, c, K, 1440]: the sum of one inverse launch per band."""
    lib, st = _ext.load(), _stream()
    dev = xm.device
    plane = _rup(c, 4) * _rup(K, 4)
    row = 2 * B * plane

    def band(b, count, scale, src, src_off):
        y = torch.empty((B, c, K, 1440), dtype=torch.float32, device=dev)
        _ext.check(lib.sfno_dft_inv_band(_fptr(src, src_off), B, c, 0, c, K, b, count, _ptr(scale), _ptr(y),
                                         dtype_code(torch.float32), st), "sfno_dft_inv_band")
        return y

    y = band(0, 241, mscale, xm, 0)
    n1 = min(M, 480) - 241
    if n1 > 0:
        # scratch row k = conj(bin 480 - k) * s[480 - k] for k = 240 - n1 .. 239; rows 0 and 240 (and the rest) are zero
        tmp = torch.zeros(240 * row, dtype=torch.float32, device=dev)
        s1 = mscale[241:241 + n1].flip(0).contiguous()       # (named: the buffers must outlive the calls that read them)
        ones = torch.ones(240, dtype=torch.float32, device=dev)
        _ext.check(lib.sfno_xm_mirror_conj(_fptr(xm, 241 * row), _fptr(tmp, (240 - n1) * row), n1, 2 * B, plane,
                                           _ptr(s1), 0, st), "sfno_xm_mirror_conj")
        y.add_(band(1, 240, ones, tmp, 0))
    if M > 480:
        s2 = mscale[480:M].clone()
        s2[0] *= 2.0                                         # the kernel counts its bin 0 once (a DC bin) ...
        if M == 721:
            s2[240] *= 0.5                                   # ... and doubles its bin 240; 720 is the Nyquist bin
        y.add_(band(2, M - 480, s2, xm, 480 * row))
    return y


def dft_fwd(x: torch.Tensor, mmax: int, mscale: torch.Tensor, round_tf32: bool = True) -> torch.Tensor:
    """x [B, C, nlat, nlon] (f32/bf16/f16, contiguous) -> XM.  XM always feeds the TF32 Legendre contraction,
    so by default its values are stored round-to-nearest TF32 (see tf32_rn in csrc/common.cuh)."""
    _need_cuda(x, "x")
    B, Cn, K, W = x.shape
    x = x.contiguous()
    xm = torch.empty(xm_elems(K, mmax, B, Cn), dtype=torch.float32, device=x.device)
    if _banded(W, mmax):
        _dft_fwd_banded(x, Cn, 0, Cn, mmax, mscale, xm, round_tf32)
        return xm
    _ext.check(_ext.load().sfno_dft_fwd(_ptr(x), dtype_code(x.dtype), B, Cn, K, W, mmax, _ptr(mscale), _ptr(xm),
                                        int(round_tf32), _stream()), "sfno_dft_fwd")
    return xm


def dft_inv(xm: torch.Tensor, B: int, Cn: int, nlat: int, nlon: int, mmax: int, mscale: torch.Tensor,
            dtype: torch.dtype, m_valid=None) -> torch.Tensor:
    _need_cuda(xm, "xm")
    if _banded(nlon, mmax if m_valid is None else min(mmax, m_valid)):
        return _dft_inv_banded(xm, B, Cn, nlat, mmax if m_valid is None else min(mmax, m_valid), mscale).to(dtype)
    y = torch.empty((B, Cn, nlat, nlon), dtype=dtype, device=xm.device)
    _ext.check(_ext.load().sfno_dft_inv(_ptr(xm), B, Cn, nlat, nlon, mmax, mmax if m_valid is None else m_valid,
                                        _ptr(mscale), _ptr(y), dtype_code(dtype), _stream()), "sfno_dft_inv")
    return y


# ------------------------------------------------------------------------------------------ Legendre
def legendre_analysis(plan: Plan, xm: torch.Tensor, B: int, Cn: int, round_tf32: bool = False) -> torch.Tensor:
    """round_tf32: the result feeds another TF32 contraction (store round-to-nearest TF32 values)."""
    sp = torch.empty(sp_elems(plan.lmax, plan.mmax, B, Cn), dtype=torch.float32, device=xm.device)
    _ext.check(_ext.load().sfno_legendre_analysis(plan.handle, _ptr(xm), B, Cn, _ptr(sp), int(round_tf32), _stream()),
               "sfno_legendre_analysis")
    return sp


def legendre_synthesis(plan: Plan, sp: torch.Tensor, B: int, Cn: int, out: torch.Tensor = None) -> torch.Tensor:
    n = xm_elems(plan.nlat, plan.mmax, B, Cn)
    if out is not None and (out.numel() != n or out.dtype != torch.float32 or not out.is_contiguous()):
        raise ValueError("legendre_synthesis: `out` must be a contiguous float32 buffer of xm_elems() floats")
    xm = out if out is not None else torch.empty(n, dtype=torch.float32, device=sp.device)
    _ext.check(_ext.load().sfno_legendre_synthesis(plan.handle, _ptr(sp), B, Cn, _ptr(xm), _stream()),
               "sfno_legendre_synthesis")
    return xm


# ------------------------------------------------------------------------------------------ channel windows
# The chain DFT -> Legendre (and Legendre -> inverse DFT) is independent per channel.  Run over WINDOWS of channels, with
# the Fourier coefficients of one window going through a small buffer that every window re-uses, the intermediate XM
# (178 MB at 240x480, 536 MB at 721x1440 for C = 384) stays in the 126 MB L2 instead of being written to HBM by one
# kernel and read back by the next: this is the fusion across the RFFT -> Legendre and Legendre -> inverse-RFFT
# boundaries of makani/models/common/spectral_convolution.py:131-141 (see include/makani_b200.h, "channel-window
# variants").  MAKANI_B200_WINDOW_MB sets the size of the window buffer; 0 (the default) runs every stage once over all
# channels.  MEASURED on B200 (round 2, profiles/r2a_window_ab.txt): the windows do take XM out of the DRAM traffic, but
# the step gets SLOWER (11.2 ms -> 13.3 ms at 120 MB, 15.3 ms at 60 MB): every extra launch costs ~15 us of ramp-up and
# tail (a window's kernels last only 20 - 35 us, a GEMM tile ~8 us), far more than the saved traffic is worth to
# kernels that are not DRAM-bound; and at 721x1440 the 167 MB Legendre table does not fit L2 next to a window, so
# every window re-reads it from DRAM.  Kept as an option (bit-identical results, tests/test_gpu_modules.py).
import os as _os

_WINDOW_MB = float(_os.environ.get("MAKANI_B200_WINDOW_MB", "0"))
_WINDOW_BUFFERS = {}


def window_channels(C: int, B: int, nlat: int, mmax: int) -> int:
    """Channels per window (a multiple of 32), or C when the chain runs in one pass."""
    if _WINDOW_MB <= 0 or C % 32 != 0:
        return C
    per_channel = mmax * 2 * B * _rup(nlat, 4) * 4
    cc = int(_WINDOW_MB * 1e6 // per_channel) // 32 * 32
    cc = max(32, cc)
    return C if cc >= C else cc


def _window_buffer(nfloats: int, device) -> torch.Tensor:
    """One persistent XM window buffer per device (grown on demand): same addresses every time, so the lines it
    occupies in L2 are simply overwritten by the next window."""
    dev = torch.device(device)
    buf = _WINDOW_BUFFERS.get(dev)
    if buf is None or buf.numel() < nfloats:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("makani_b200: run one eager warm-up step before CUDA-graph capture (window buffer)")
        buf = torch.empty(nfloats, dtype=torch.float32, device=dev)
        _WINDOW_BUFFERS[dev] = buf
    return buf[:nfloats]


def sht_to_sp(plan: Plan, x: torch.Tensor, mscale: torch.Tensor, round_tf32: bool = False, out: torch.Tensor = None,
              accumulate: bool = False) -> torch.Tensor:
    """x [B, C, nlat, nlon] -> SP: longitude DFT (K1) + Legendre analysis (K2), window by window.
    accumulate: add to `out` (the sum of the gradients of two consumers of one SP tensor)."""
    _need_cuda(x, "x")
    B, Cn, K, W = x.shape
    x = x.contiguous()
    M = plan.mmax
    sp = out if out is not None else torch.empty(sp_elems(plan.lmax, M, B, Cn), dtype=torch.float32, device=x.device)
    cc = window_channels(Cn, B, K, M)
    if cc >= Cn:
        xm = dft_fwd(x, M, mscale)
        _ext.check(_ext.load().sfno_legendre_analysis_win(plan.handle, _ptr(xm), B, Cn, 0, Cn, _ptr(sp), int(round_tf32),
                                                          int(accumulate), _stream()), "sfno_legendre_analysis_win")
        return sp
    lib, st, dt = _ext.load(), _stream(), dtype_code(x.dtype)
    xm = _window_buffer(xm_elems(K, M, B, cc), x.device)
    for c0 in range(0, Cn, cc):
        c = min(cc, Cn - c0)
        if _banded(W, M):
            _dft_fwd_banded(x, Cn, c0, c, M, mscale, xm, True)
        else:
            _ext.check(lib.sfno_dft_fwd_win(_ptr(x), dt, B, Cn, c0, c, K, W, M, _ptr(mscale), _ptr(xm), 1, st),
                       "sfno_dft_fwd_win")
        _ext.check(lib.sfno_legendre_analysis_win(plan.handle, _ptr(xm), B, Cn, c0, c, _ptr(sp), int(round_tf32),
                                                  int(accumulate), st), "sfno_legendre_analysis_win")
    return sp


def sp_to_field(plan: Plan, sp: torch.Tensor, B: int, Cn: int, nlon: int, mscale: torch.Tensor, dtype: torch.dtype,
                m_valid=None, dc_bias: torch.Tensor = None) -> torch.Tensor:
    """SP -> y [B, C, nlat, nlon]: Legendre synthesis (K4) + inverse longitude DFT (K5), window by window.
    dc_bias: fp32 [C], added to every point of channel c in fp32 (as the m = 0 coefficient; needs mscale[0] == 1)."""
    K, M = plan.nlat, plan.mmax
    mv = M if m_valid is None else m_valid
    cc = window_channels(Cn, B, K, M)
    if cc >= Cn:
        xm = legendre_synthesis(plan, sp, B, Cn)
        if dc_bias is not None:
            _ext.check(_ext.load().sfno_xm_add_dc(_ptr(xm), B, Cn, K, _ptr(dc_bias), _stream()), "sfno_xm_add_dc")
        return dft_inv(xm, B, Cn, K, nlon, M, mscale, dtype, m_valid)
    lib, st, dt = _ext.load(), _stream(), dtype_code(dtype)
    y = torch.empty((B, Cn, K, nlon), dtype=dtype, device=sp.device)
    xm = _window_buffer(xm_elems(K, M, B, cc), sp.device)
    for c0 in range(0, Cn, cc):
        c = min(cc, Cn - c0)
        _ext.check(lib.sfno_legendre_synthesis_win(plan.handle, _ptr(sp), B, Cn, c0, c, _ptr(xm), st),
                   "sfno_legendre_synthesis_win")
        if dc_bias is not None:
            _ext.check(lib.sfno_xm_add_dc(_ptr(xm), B, c, K, C.c_void_p(dc_bias.data_ptr() + 4 * c0), st),
                       "sfno_xm_add_dc")
        if _banded(nlon, min(M, mv)):
            y[:, c0:c0 + c].copy_(_dft_inv_banded(xm, B, c, K, min(M, mv), mscale))
        else:
            _ext.check(lib.sfno_dft_inv_win(_ptr(xm), B, Cn, c0, c, K, nlon, M, mv, _ptr(mscale), _ptr(y), dt, st),
                       "sfno_dft_inv_win")
    return y


# ------------------------------------------------------------------------------------------ layout
def sp_to_c64(sp, B, Cn, lmax, mmax, l_offset=0, m_offset=0) -> torch.Tensor:
    out = torch.empty((B, Cn, lmax, mmax), dtype=torch.complex64, device=sp.device)
    _ext.check(_ext.load().sfno_sp_to_c64(_ptr(sp), B, Cn, lmax, mmax, l_offset, m_offset, _ptr(out), _stream()),
               "sfno_sp_to_c64")
    return out


def c64_to_sp(c: torch.Tensor, l_offset: int = 0, m_offset: int = 0, round_tf32: bool = True) -> torch.Tensor:
    """complex64 [B, C, L, M] -> SP.  Entries with m + m_offset > l + l_offset are taken as zero (not read).
    round_tf32: the consumer is a TF32 tensor-core contraction (store round-to-nearest TF32 values)."""
    _need_cuda(c, "coefficients")
    B, Cn, L, M = c.shape
    c = c.contiguous()
    if c.dtype != torch.complex64:
        raise TypeError(f"makani_b200: spectral coefficients must be complex64, got {c.dtype}")
    sp = torch.empty(sp_elems(L, M, B, Cn), dtype=torch.float32, device=c.device)
    _ext.check(_ext.load().sfno_c64_to_sp(_ptr(c), B, Cn, L, M, l_offset, m_offset, int(round_tf32), _ptr(sp),
                                          _stream()),
               "sfno_c64_to_sp")
    return sp


def xm_to_c64(xm, B, Cn, nlat, mmax) -> torch.Tensor:
    out = torch.empty((B, Cn, nlat, mmax), dtype=torch.complex64, device=xm.device)
    _ext.check(_ext.load().sfno_xm_to_c64(_ptr(xm), B, Cn, nlat, mmax, _ptr(out), _stream()), "sfno_xm_to_c64")
    return out


def c64_to_xm(c: torch.Tensor) -> torch.Tensor:
    _need_cuda(c, "coefficients")
    B, Cn, K, M = c.shape
    c = c.contiguous()
    if c.dtype != torch.complex64:
        raise TypeError(f"makani_b200: Fourier coefficients must be complex64, got {c.dtype}")
    xm = torch.empty(xm_elems(K, M, B, Cn), dtype=torch.float32, device=c.device)
    _ext.check(_ext.load().sfno_c64_to_xm(_ptr(c), B, Cn, K, M, _ptr(xm), _stream()), "sfno_c64_to_xm")
    return xm


def add_bias_(y: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    """In-place y[b,c,h,w] += bias, bias [1,C,1,1] or [1,C,H,W] (fp32)."""
    B, Cn, H, W = y.shape
    kind = 0 if bias.numel() == Cn else 1
    bias = bias.detach().to(torch.float32).contiguous()
    _ext.check(_ext.load().sfno_add_bias(_ptr(y), dtype_code(y.dtype), _ptr(bias), kind, B, Cn, H, W, _stream()),
               "sfno_add_bias")
    return y


def sp_band_add_(dst: torch.Tensor, src: torch.Tensor, B: int, Cn: int, lmax: int, mmax: int, l_offset: int = 0):
    """dst += src for two SP tensors of the same shape (values outside the defined band are never read by anyone)."""
    n = sp_elems(lmax, mmax, B, Cn)
    _ext.check(_ext.load().sfno_add_inplace(_ptr(dst), _ptr(src), n, _stream()), "sfno_add_inplace")
    return dst


def sp_degree_power(sp: torch.Tensor, B: int, Cn: int, lmax: int, mmax: int, l_offset=0, m_offset=0) -> torch.Tensor:
    """SP -> fp32 [B, C, lmax]: sum over m <= l of |coefficient|^2, counted once at m = 0 and twice elsewhere
    (makani/utils/losses.py:299-303)."""
    out = torch.empty((B, Cn, lmax), dtype=torch.float32, device=sp.device)
    _ext.check(_ext.load().sfno_sp_degree_power(_ptr(sp), B, Cn, lmax, mmax, l_offset, m_offset, _ptr(out), _stream()),
               "sfno_sp_degree_power")
    return out


def sp_degree_power_bwd(sp: torch.Tensor, g: torch.Tensor, B: int, Cn: int, lmax: int, mmax: int, l_offset=0,
                        m_offset=0, round_tf32: bool = True) -> torch.Tensor:
    """Gradient of sp_degree_power with respect to SP, as an SP tensor (zeros above the diagonal and in pad channels)."""
    g = g.to(torch.float32).contiguous()
    if tuple(g.shape) != (B, Cn, lmax):
        raise ValueError(f"makani_b200: expected a gradient of shape {(B, Cn, lmax)}, got {tuple(g.shape)}")
    gsp = torch.empty_like(sp)
    _ext.check(_ext.load().sfno_sp_degree_power_bwd(_ptr(sp), _ptr(g), B, Cn, lmax, mmax, l_offset, m_offset,
                                                    int(round_tf32), _ptr(gsp), _stream()), "sfno_sp_degree_power_bwd")
    return gsp


# ------------------------------------------------------------------------------------------ dhconv
def pack_weight(w: torch.Tensor, round_tf32: bool = True) -> torch.Tensor:
    """complex64 [Cin, Cout, L] -> WP.  round_tf32=False keeps the fp32 values (packed master copy)."""
    _need_cuda(w, "weight")
    Cin, Cout, L = w.shape
    w = w.detach().contiguous()
    wp = torch.empty(wp_elems(L, Cin, Cout), dtype=torch.float32, device=w.device)
    _ext.check(_ext.load().sfno_dhconv_pack_weight_ex(_ptr(w), Cin, Cout, L, _ptr(wp), int(round_tf32), _stream()),
               "sfno_dhconv_pack_weight")
    return wp


def unpack_wgrad(gwp: torch.Tensor, Cin, Cout, L) -> torch.Tensor:
    gw = torch.empty((Cin, Cout, L), dtype=torch.complex64, device=gwp.device)
    _ext.check(_ext.load().sfno_dhconv_unpack_wgrad(_ptr(gwp), Cin, Cout, L, _ptr(gw), _stream()),
               "sfno_dhconv_unpack_wgrad")
    return gw


def dhconv_fwd(sp_in, wp, B, Cin, Cout, lmax, mmax, l_offset=0, m_offset=0, round_tf32=False) -> torch.Tensor:
    sp_out = torch.empty(sp_elems(lmax, mmax, B, Cout), dtype=torch.float32, device=sp_in.device)
    _ext.check(_ext.load().sfno_dhconv_fwd(_ptr(sp_in), _ptr(wp), _ptr(sp_out), B, Cin, Cout, lmax, mmax,
                                           l_offset, m_offset, int(round_tf32), _stream()), "sfno_dhconv_fwd")
    return sp_out


def dhconv_bwd_x(gsp_out, wp, B, Cin, Cout, lmax, mmax, l_offset=0, m_offset=0, round_tf32=False) -> torch.Tensor:
    gsp_in = torch.empty(sp_elems(lmax, mmax, B, Cin), dtype=torch.float32, device=gsp_out.device)
    _ext.check(_ext.load().sfno_dhconv_bwd_x(_ptr(gsp_out), _ptr(wp), _ptr(gsp_in), B, Cin, Cout, lmax, mmax,
                                             l_offset, m_offset, int(round_tf32), _stream()), "sfno_dhconv_bwd_x")
    return gsp_in


def dhconv_bwd_w(sp_in, gsp_out, B, Cin, Cout, lmax, mmax, l_offset=0, m_offset=0, out=None) -> torch.Tensor:
    n = wp_elems(lmax, Cin, Cout)
    if out is not None and (out.numel() != n or out.dtype != torch.float32 or not out.is_contiguous()):
        raise ValueError("dhconv_bwd_w: `out` must be a contiguous float32 buffer of wp_elems() floats")
    gwp = out if out is not None else torch.empty(n, dtype=torch.float32, device=sp_in.device)
    _ext.check(_ext.load().sfno_dhconv_bwd_w(_ptr(sp_in), _ptr(gsp_out), _ptr(gwp), B, Cin, Cout, lmax, mmax,
                                             l_offset, m_offset, _stream()), "sfno_dhconv_bwd_w")
    return gwp


# ------------------------------------------------------------------------------------------ diagonal
def diag_fwd(sp_in, w, B, Cin, Cout, lmax, mmax, l_offset=0, m_offset=0) -> torch.Tensor:
    w = w.detach().contiguous()
    sp_out = torch.empty(sp_elems(lmax, mmax, B, Cout), dtype=torch.float32, device=sp_in.device)
    _ext.check(_ext.load().sfno_diag_contract(0, _ptr(sp_in), _ptr(w), _ptr(sp_out), B, Cin, Cout, lmax, mmax,
                                              l_offset, m_offset, _stream()), "sfno_diag_contract(fwd)")
    return sp_out


def diag_bwd_x(gsp_out, w, B, Cin, Cout, lmax, mmax, l_offset=0, m_offset=0) -> torch.Tensor:
    w = w.detach().contiguous()
    gsp_in = torch.empty(sp_elems(lmax, mmax, B, Cin), dtype=torch.float32, device=gsp_out.device)
    _ext.check(_ext.load().sfno_diag_contract(1, _ptr(gsp_out), _ptr(w), _ptr(gsp_in), B, Cin, Cout, lmax, mmax,
                                              l_offset, m_offset, _stream()), "sfno_diag_contract(bwd_x)")
    return gsp_in


def diag_bwd_w(sp_in, gsp_out, B, Cin, Cout, lmax, mmax, l_offset=0, m_offset=0) -> torch.Tensor:
    gw = torch.empty((Cin, Cout, lmax, mmax), dtype=torch.complex64, device=sp_in.device)
    _ext.check(_ext.load().sfno_diag_contract(2, _ptr(sp_in), _ptr(gsp_out), _ptr(gw), B, Cin, Cout, lmax, mmax,
                                              l_offset, m_offset, _stream()), "sfno_diag_contract(bwd_w)")
    return gw


# ------------------------------------------------------------------------------------------ composite calls
def _workspace(plan: Plan, plan_b, op: int, B: int, Cn: int, device) -> torch.Tensor:
    n = int(_ext.load().sfno_workspace_bytes(plan.handle, None if plan_b is None else plan_b.handle, op, B, Cn))
    if n == 0:
        raise RuntimeError("makani_b200: sfno_workspace_bytes returned 0 (bad plan / sizes)")
    return torch.empty(n, dtype=torch.uint8, device=device)


def sht_fwd(plan: Plan, x: torch.Tensor) -> torch.Tensor:
    """One-call RealSHT.forward (sfno_sht_fwd): x [B, C, nlat, nlon] -> complex64 [B, C, lmax, mmax]."""
    _need_cuda(x, "x")
    B, Cn, K, W = x.shape
    x = x.contiguous()
    ws = _workspace(plan, None, _ext.OP_SHT_FWD, B, Cn, x.device)
    out = torch.empty((B, Cn, plan.lmax, plan.mmax), dtype=torch.complex64, device=x.device)
    _ext.check(_ext.load().sfno_sht_fwd(plan.handle, _ptr(x), dtype_code(x.dtype), B, Cn, _ptr(out), _ptr(ws),
                                        ws.numel(), _stream()), "sfno_sht_fwd")
    return out


def sht_inv(plan: Plan, c: torch.Tensor, dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """One-call InverseRealSHT.forward (sfno_sht_inv): complex64 [B, C, lmax, mmax] -> [B, C, nlat, nlon]."""
    _need_cuda(c, "coefficients")
    B, Cn = c.shape[0], c.shape[1]
    c = c.contiguous()
    ws = _workspace(plan, None, _ext.OP_SHT_INV, B, Cn, c.device)
    y = torch.empty((B, Cn, plan.nlat, plan.nlon), dtype=dtype, device=c.device)
    _ext.check(_ext.load().sfno_sht_inv(plan.handle, _ptr(c), B, Cn, _ptr(y), dtype_code(dtype), _ptr(ws),
                                        ws.numel(), _stream()), "sfno_sht_inv")
    return y


def block_fwd(plan_fwd: Plan, plan_inv: Plan, x: torch.Tensor, wp: torch.Tensor, Cout: int) -> torch.Tensor:
    """One-call SpectralConv.forward for dhconv without bias (sfno_block_fwd): y = isht(dhconv(sht(x), w))."""
    _need_cuda(x, "x")
    B, Cin, K, W = x.shape
    x = x.contiguous()
    ws = _workspace(plan_fwd, plan_inv, _ext.OP_BLOCK_FWD, B, max(Cin, Cout), x.device)
    y = torch.empty((B, Cout, plan_inv.nlat, plan_inv.nlon), dtype=x.dtype, device=x.device)
    _ext.check(_ext.load().sfno_block_fwd(plan_fwd.handle, plan_inv.handle, _ptr(x), dtype_code(x.dtype), B, Cin, Cout,
                                          _ptr(wp), _ptr(y), _ptr(ws), ws.numel(), _stream()), "sfno_block_fwd")
    return y
