"""CPU tests of the host-side logic: fp64 precompute vs the oracle, the C-ABI library (loads, exports every
symbol the header declares - no compute calls without a GPU), module construction / API surface, the host
emulation of the device FFT lane code, and loud failure on CPU tensors."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest
import torch

from oracle import sht_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_precompute_matches_oracle():
    from makani_b200 import legendre, quadrature
    for grid, nlat in [("legendre-gauss", 24), ("equiangular", 33), ("lobatto", 17)]:
        cost, w, _ = O.grid_nodes_weights(nlat, grid)
        fn = {"legendre-gauss": quadrature.legendre_gauss_weights, "equiangular": quadrature.clenshaw_curtiss_weights,
              "lobatto": quadrature.lobatto_weights}[grid]
        c2, w2 = fn(nlat, -1, 1)
        np.testing.assert_allclose(c2, cost, rtol=0, atol=1e-15)
        np.testing.assert_allclose(w2, w, rtol=0, atol=1e-15)
        tq = np.flip(np.arccos(cost))
        for inverse in (False, True):
            a = O.precompute_legpoly(13, 20, tq, inverse=inverse)
            b = legendre.legpoly(13, 20, tq, inverse=inverse)
            np.testing.assert_array_equal(a, b)
    # 721-point table used by the full-resolution transforms: finite, bounded, triangular
    cost, _ = quadrature.clenshaw_curtiss_weights(721)
    tab = legendre.legpoly(41, 40, np.flip(np.arccos(cost)))
    assert np.isfinite(tab).all() and np.abs(tab).max() < 10
    assert (tab[5, :5] == 0).all()
    t, wq = quadrature.precompute_latitudes(16, "legendre-gauss")
    assert t[0] < t[-1] and abs(wq.sum() - 2.0) < 1e-12
    n, wf = quadrature.fejer2_weights(17)
    assert abs(wf.sum() - 2.0) < 1e-12


def test_library_exports_every_declared_symbol(built_lib):
    header = open(os.path.join(ROOT, "include", "makani_b200.h")).read()
    declared = set(re.findall(r"\b(sfno_[a-z0-9_]+)\s*\(", header))
    declared.discard("sfno_plan")
    assert len(declared) >= 20
    lib = ctypes.CDLL(built_lib)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/makani_b200.h but not exported"
    from makani_b200 import _ext
    assert set(_ext.EXPORTED_SYMBOLS) == declared
    assert _ext.load().sfno_abi_version() == _ext.ABI_VERSION == 3
    assert _ext.launch_count() == 0
    # size helpers are pure host arithmetic
    assert _ext.load().sfno_xm_elems(721, 241, 1, 384) == 241 * 768 * 724
    assert _ext.load().sfno_sp_elems(240, 241, 2, 6) == 240 * 241 * 2 * 2 * 8
    assert _ext.load().sfno_wp_elems(240, 384, 384) == 240 * 2 * 384 * 384


def test_library_has_tcgen05_and_tma_sass(built_lib):
    out = subprocess.run(["cuobjdump", "-sass", built_lib], capture_output=True, text=True).stdout
    assert "UTCHMMA" in out or "UTCQMMA" in out, "no tcgen05.mma in SASS"
    assert "UTMALDG" in out, "no TMA loads in SASS"
    assert "LDTM" in out, "no tcgen05.ld in SASS"


@pytest.mark.parametrize("name,token", [("fft_lane_host_test", "FFT_LANE_OK"), ("fft_pass_host_test", "FFT_PASS_OK")])
def test_fft_host_emulation(name, token):
    """The per-lane FFT code of the DFT kernels (csrc/fft_lane.cuh: 16-lane shuffle FFT; csrc/fft_pass.cuh: two-pass
    16 x 15 FFT) compiled for the host, one thread per lane, against a direct double-precision DFT."""
    exe = os.path.join(ROOT, "tests", "cpp", name + ".bin")
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", os.path.join(ROOT, "tests", "cpp", name + ".cpp"),
                    "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and token in r.stdout, r.stdout + r.stderr


def test_module_surface_and_cpu_refusal(built_lib):
    import makani_b200 as mk
    sht = mk.RealSHT(32, 64, grid="legendre-gauss").float()
    isht = mk.InverseRealSHT(32, 64, grid="legendre-gauss").float()
    assert (sht.nlat, sht.nlon, sht.lmax, sht.mmax, sht.grid) == (32, 64, 32, 33, "legendre-gauss")
    assert mk.RealSHT(33, 64).lmax == 32                     # lobatto default lmax = nlat - 1
    assert sht.weights.dtype == torch.float32 and tuple(sht.weights.shape) == (33, 32, 32)
    assert tuple(isht.pct.shape) == (33, 32, 32)
    assert list(sht.state_dict()) == [] and list(isht.state_dict()) == []
    ref = O.RealSHT(32, 64, grid="legendre-gauss")
    np.testing.assert_allclose(sht.weights.double().numpy(), ref.weights_f64, rtol=0, atol=1e-7)
    conv = mk.SpectralConv(sht, isht, 4, 4, operator_type="dhconv")
    assert not conv.scale_residual
    conv2 = mk.SpectralConv(mk.RealSHT(48, 96, lmax=32, mmax=33, grid="equiangular"), isht, 4, 4, operator_type="dhconv")
    assert conv2.scale_residual
    for call in (lambda: sht(torch.randn(1, 32, 64)), lambda: isht(torch.randn(1, 32, 33, dtype=torch.complex64)),
                 lambda: conv(torch.randn(1, 4, 32, 64))):
        with pytest.raises(RuntimeError, match="no CPU path"):
            call()
    with pytest.raises(NotImplementedError):
        mk.SpectralConv(sht, isht, 4, 4, operator_type="dhconv", separable=True)


def test_compat_shim(built_lib):
    import sys
    import makani_b200 as mk
    saved = {k: v for k, v in sys.modules.items() if k.startswith("torch_harmonics")}
    try:
        for k in saved:
            del sys.modules[k]
        mk.install_torch_harmonics_shim()
        import torch_harmonics as th
        import torch_harmonics.distributed as thd
        from torch_harmonics.quadrature import legendre_gauss_weights, clenshaw_curtiss_weights
        assert th.RealSHT is mk.RealSHT and th.InverseRealSHT is mk.InverseRealSHT
        assert thd.DistributedRealSHT is mk.distributed.DistributedRealSHT
        assert callable(thd.init) and callable(thd.distributed_transpose_polar.apply)
        x, w = legendre_gauss_weights(8, -1, 1)
        assert abs(w.sum() - 2) < 1e-14
    finally:
        for k in [k for k in sys.modules if k.startswith("torch_harmonics")]:
            del sys.modules[k]
        sys.modules.update(saved)
