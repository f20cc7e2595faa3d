"""TEST INFRASTRUCTURE: a host emulation of the four ``sfno_instnorm_*`` entry points (include/makani_b200.h) on CPU
pointers, fp32 only, written from the header's contract with numpy.  It lets the CPU suite drive the HOST logic of
``makani_b200.instance_norm`` - the sharded moment merge, the backward all-reduce, which gradients are partial - under
gloo, where the CUDA kernels cannot run.  It is not a fallback: nothing in the product imports it."""
import ctypes as C
import math

import numpy as np

SEG = 8


def _arr(ptr, n, dtype=np.float32):
    addr = ptr.value if isinstance(ptr, C.c_void_p) else ptr
    if not addr:
        return None
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(addr)
    return np.frombuffer(buf, dtype=dtype, count=n)


def _gelu(z):
    erf = np.vectorize(math.erf)
    return 0.5 * z * (1 + erf(z / math.sqrt(2)))


def _gelu_grad(z):
    erf = np.vectorize(math.erf)
    return 0.5 * (1 + erf(z / math.sqrt(2))) + z * np.exp(-0.5 * z * z) / math.sqrt(2 * math.pi)


class FakeLib:
    def __init__(self):
        self.calls = []

    def sfno_last_error(self):
        return b"fake"

    def sfno_instnorm_stats(self, x, dtype, planes, hw, eps, ws, mean_var, stats, stream):
        assert dtype == 0
        self.calls.append("stats")
        xa = _arr(x, planes * hw).reshape(planes, hw).astype(np.float64)
        mean, var = xa.mean(axis=1), xa.var(axis=1)
        mv, st = _arr(mean_var, planes * 2), _arr(stats, planes * 2)
        if mv is not None:
            mv.reshape(planes, 2)[:] = np.stack([mean, var], axis=1)
        if st is not None:
            st.reshape(planes, 2)[:] = np.stack([mean, 1 / np.sqrt(var + eps)], axis=1)
        return 0

    def _common(self, x, planes, Cn, hw, stats, gamma, beta):
        xa = _arr(x, planes * hw).reshape(planes, hw).astype(np.float64)
        st = _arr(stats, planes * 2).reshape(planes, 2).astype(np.float64)
        c = np.arange(planes) % Cn
        ga = _arr(gamma, Cn)
        be = _arr(beta, Cn)
        ga = np.ones(planes) if ga is None else ga.astype(np.float64)[c]
        be = np.zeros(planes) if be is None else be.astype(np.float64)[c]
        xh = (xa - st[:, :1]) * st[:, 1:]
        return xh, st, ga[:, None], be[:, None]

    def sfno_instnorm_apply(self, x, skip, dtype, planes, Cn, hw, stats, gamma, beta, act, out, stream):
        self.calls.append("apply")
        xh, st, ga, be = self._common(x, planes, Cn, hw, stats, gamma, beta)
        z = ga * xh + be
        sk = _arr(skip, planes * hw)
        if sk is not None:
            z = z + sk.reshape(planes, hw)
        _arr(out, planes * hw).reshape(planes, hw)[:] = _gelu(z) if act == 1 else z
        return 0

    def sfno_instnorm_bwd_stats(self, g, x, dtype, planes, Cn, hw, stats, gamma, beta, act, partial, stream):
        self.calls.append("bwd_stats")
        xh, st, ga, be = self._common(x, planes, Cn, hw, stats, gamma, beta)
        dz = _arr(g, planes * hw).reshape(planes, hw).astype(np.float64)
        if act == 1:
            dz = dz * _gelu_grad(ga * xh + be)
        part = _arr(partial, planes * SEG * 2).reshape(planes, SEG, 2)
        part[:] = 0
        part[:, 0, 0] = dz.sum(axis=1)            # all of it in segment 0: the caller sums the segments
        part[:, 0, 1] = (dz * xh).sum(axis=1)
        return 0

    def sfno_instnorm_bwd_apply(self, g, x, dtype, planes, Cn, hw, hw_total, stats, gamma, beta, bsum, act, dx, stream):
        self.calls.append("bwd_apply")
        xh, st, ga, be = self._common(x, planes, Cn, hw, stats, gamma, beta)
        dz = _arr(g, planes * hw).reshape(planes, hw).astype(np.float64)
        if act == 1:
            dz = dz * _gelu_grad(ga * xh + be)
        bs = _arr(bsum, planes * 2).reshape(planes, 2).astype(np.float64)
        _arr(dx, planes * hw).reshape(planes, hw)[:] = ga * st[:, 1:] * (dz - (bs[:, :1] + xh * bs[:, 1:]) / hw_total)
        return 0


def install(monkeypatch_or_none=None):
    """Point makani_b200.instance_norm at the emulation (and lift its CUDA-only guard).  Returns the FakeLib."""
    import makani_b200.instance_norm as inorm
    import makani_b200._ext as ext
    fake = FakeLib()
    ext._real_load = getattr(ext, "_real_load", ext.load)
    ext.load = lambda: fake
    ext.check = lambda status, what: (_ for _ in ()).throw(RuntimeError(what)) if status != 0 else None
    inorm._need_cuda = lambda t, name: None
    inorm._stream = lambda: C.c_void_p(0)
    return fake
