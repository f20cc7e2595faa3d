"""Kernel-level timing for bench.py: every C-ABI call of libmakani_b200.so that launches a kernel of the hot path is
bracketed by CUDA events ON THE LAUNCHING STREAM (torch's current stream - the one the wrappers in makani_b200/ops.py
pass to the library), for one profiling pass outside the timed region.  The ctypes function objects of the loaded
library are wrapped in place, so the numbers cover exactly what the modules call, window by window."""
import torch

from . import workload as W


def _esize(code):
    return 4 if code == 0 else 2


# name -> (family, describe(args) -> shape dict).  Argument positions follow include/makani_b200.h.
def _d_dft_fwd(a):       # x, dtype, B, C, nlat, nlon, mmax, mscale, xm, round, stream
    return dict(B=a[2], C=a[3], K=a[4], W=a[5], M=a[6], esize=_esize(a[1]))


def _d_dft_inv(a):       # xm, B, C, nlat, nlon, mmax, m_valid, mscale, y, dtype, stream
    return dict(B=a[1], C=a[2], K=a[3], W=a[4], M=a[5], esize=_esize(a[9]))


def _d_dft_fwd_win(a):   # x, dtype, B, Ctot, c0, C, nlat, nlon, mmax, ...
    return dict(B=a[2], C=a[5], K=a[6], W=a[7], M=a[8], esize=_esize(a[1]))


def _d_dft_inv_win(a):   # xm, B, Ctot, c0, C, nlat, nlon, mmax, m_valid, mscale, y, dtype, stream
    return dict(B=a[1], C=a[4], K=a[5], W=a[6], M=a[7], esize=_esize(a[11]))


class _PlanDims:
    """plan handle -> (nlat, lmax, mmax); filled by the caller from the modules' plans."""
    dims = {}


def _plan(a0):
    return _PlanDims.dims.get(getattr(a0, "value", a0), (0, 0, 0))


def _d_leg(a):           # plan, in, B, C, out, ...
    K, L, M = _plan(a[0])
    return dict(B=a[2], C=a[3], K=K, L=L, M=M, c0=0)


def _d_leg_win(a):       # plan, in, B, Ctot, c0, C, out, ...
    K, L, M = _plan(a[0])
    return dict(B=a[2], C=a[5], K=K, L=L, M=M, c0=a[4])


def _d_dh(a):            # in, w, out, B, Cin, Cout, lmax, mmax, ...
    return dict(B=a[3], Cin=a[4], Cout=a[5], L=a[6], M=a[7])


def _d_pack(a):          # w, Cin, Cout, L, wp, stream
    return dict(Cin=a[1], Cout=a[2], L=a[3])


def _d_gather(a):        # slices, nslices, slice_floats, Cin, Cout, L, gw, stream
    return dict(Cin=a[3], Cout=a[4], L=a[5], peers=a[1])


def _d_avg(a):           # dst, srcs, nsrc, n, scale, stream
    return dict(n=a[3], nsrc=a[2])


CALLS = {
    "sfno_dft_fwd": ("dft_fwd", _d_dft_fwd), "sfno_dft_fwd_win": ("dft_fwd", _d_dft_fwd_win),
    "sfno_dft_inv": ("dft_inv", _d_dft_inv), "sfno_dft_inv_win": ("dft_inv", _d_dft_inv_win),
    "sfno_legendre_analysis": ("legendre_analysis", _d_leg), "sfno_legendre_analysis_win": ("legendre_analysis", _d_leg_win),
    "sfno_legendre_synthesis": ("legendre_synthesis", _d_leg), "sfno_legendre_synthesis_win": ("legendre_synthesis", _d_leg_win),
    "sfno_dhconv_fwd": ("dhconv_fwd", _d_dh), "sfno_dhconv_bwd_x": ("dhconv_bwd_x", _d_dh),
    "sfno_dhconv_bwd_w": ("dhconv_bwd_w", _d_dh),
    "sfno_dhconv_pack_weight": ("pack_weight", _d_pack), "sfno_dhconv_pack_weight_ex": ("pack_weight", _d_pack),
    "sfno_dhconv_unpack_wgrad": ("unpack_wgrad", _d_pack),
    "sfno_dhconv_unpack_wgrad_gather": ("unpack_wgrad_gather", _d_gather), "sfno_slices_average": ("grad_slice_average", _d_avg),
}


class KernelProfiler:
    def __init__(self, plans=()):
        from makani_b200 import _ext
        self.lib = _ext.load()
        self.records = []
        self.saved = {}
        for p in plans:
            _PlanDims.dims[p.handle.value] = (p.nlat, p.lmax, p.mmax)

    def _wrap(self, family, fn, describe):
        def inner(*args):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(*args)
            e1.record()
            self.records.append((family, describe(args), e0, e1))
            return rc
        return inner

    def __enter__(self):
        for name, (family, describe) in CALLS.items():
            fn = getattr(self.lib, name)
            self.saved[name] = fn
            setattr(self.lib, name, self._wrap(family, fn, describe))
        return self

    def __exit__(self, *exc):
        for name, fn in self.saved.items():
            setattr(self.lib, name, fn)

    def summary(self):
        """Rows per (kernel family, grid): the channel windows of one transform are summed into one row (their bytes
        and their time add up to the whole-tensor figures), `launches` counts kernel launches."""
        torch.cuda.synchronize()
        agg = {}
        for family, a, e0, e1 in self.records:
            key_shape = {k: v for k, v in a.items() if k not in ("C", "c0")}
            key = (family, tuple(sorted(key_shape.items())))
            t = agg.setdefault(key, dict(name=family, shape=key_shape, ms=0.0, n=0, bytes=0, chan=0))
            t["ms"] += e0.elapsed_time(e1)
            t["n"] += 1
            t["bytes"] += W.kernel_bytes(family, a)
            t["chan"] += a.get("C", 0)
        rows = []
        for t in agg.values():
            rows.append(dict(kernel=t["name"], shape=t["shape"], launches=t["n"], total_ms=t["ms"],
                             alg_bytes_total=t["bytes"], channels_total=t["chan"],
                             gbs=t["bytes"] / (t["ms"] * 1e-3) / 1e9 if t["ms"] > 0 else 0.0))
        rows.sort(key=lambda r: -r["total_ms"])
        return rows
