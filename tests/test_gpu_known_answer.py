"""Known-answer tests of the CUDA transforms against a THIRD-PARTY definition of the spherical harmonics
(scipy.special.sph_harm_y), independent of oracle/sht_oracle.py.

The reference's RealSHT (torch-harmonics, norm="ortho", Condon-Shortley phase) is the orthonormal expansion
    f(theta, phi) = sum_l [ c_l0 Y_l^0 + sum_{m>0} 2 Re( c_lm Y_l^m ) ],
so on the SFNO grids (240x480 Gauss-Legendre and 721x1440 equiangular / Clenshaw-Curtis, L = 240, M = 241 - the
transforms constructed at makani/models/networks/sfnonet.py:536-539):

    RealSHT( Y_l^0 )        = e_(l,0)          RealSHT( 2 Re Y_l^m ) = e_(l,m)        RealSHT( 2 Im Y_l^m ) = -i e_(l,m)
    InverseRealSHT( e_(l,m) ) = Y_l^0  (m = 0)  or  2 Re Y_l^m  (m > 0)

for (l, m) in {(0,0), (7,3), (239,0), (239,1), (239,239)}.  Tolerance: TOL_T = 1e-3 relative L2 (one TF32
contraction), the same as the oracle-based module tests.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL_T = 1e-3
DEV = "cuda:0"
L, M = 240, 241
MODES = [(0, 0), (7, 3), (239, 0), (239, 1), (239, 239)]
GRIDS = [(240, 480, "legendre-gauss"), (721, 1440, "equiangular")]


@pytest.fixture(scope="module")
def mk(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import makani_b200
    return makani_b200


def _fields(nlat, nlon, grid):
    """Stack of real fields [n, nlat, nlon] (fp64) and the coefficient each one must produce."""
    from scipy.special import sph_harm_y
    from makani_b200.quadrature import clenshaw_curtiss_weights, legendre_gauss_weights
    cost, _ = (legendre_gauss_weights if grid == "legendre-gauss" else clenshaw_curtiss_weights)(nlat, -1, 1)
    theta = np.flip(np.arccos(cost))                       # colatitudes, north -> south
    phi = np.linspace(0, 2 * np.pi, nlon, endpoint=False)
    fields, expect = [], []
    for (l, m) in MODES:
        Y = sph_harm_y(l, m, theta[:, None], phi[None, :])
        if m == 0:
            fields.append(Y.real)
            expect.append((l, m, 1.0 + 0.0j))
        else:
            fields.append(2.0 * Y.real)
            expect.append((l, m, 1.0 + 0.0j))
            fields.append(2.0 * Y.imag)
            expect.append((l, m, -1.0j))
    return np.stack(fields), expect


def _rel(a, b):
    return float((a - b).norm() / b.norm())


@pytest.mark.parametrize("nlat,nlon,grid", GRIDS)
def test_real_sht_of_spherical_harmonics_is_a_unit_coefficient(mk, nlat, nlon, grid):
    f, expect = _fields(nlat, nlon, grid)
    sht = mk.RealSHT(nlat, nlon, lmax=L, mmax=M, grid=grid).to(DEV)
    x = torch.from_numpy(f).float().unsqueeze(0).to(DEV)            # [1, n, nlat, nlon]
    c = sht(x)[0].cpu().to(torch.complex128)
    for i, (l, m, v) in enumerate(expect):
        e = torch.zeros(L, M, dtype=torch.complex128)
        e[l, m] = v
        err = _rel(c[i], e)
        assert err < TOL_T, (grid, l, m, v, err)


@pytest.mark.parametrize("nlat,nlon,grid", GRIDS)
def test_inverse_sht_of_a_unit_coefficient_is_the_spherical_harmonic(mk, nlat, nlon, grid):
    f, expect = _fields(nlat, nlon, grid)
    isht = mk.InverseRealSHT(nlat, nlon, lmax=L, mmax=M, grid=grid).to(DEV)
    c = torch.zeros(1, len(expect), L, M, dtype=torch.complex64)
    for i, (l, m, v) in enumerate(expect):
        c[0, i, l, m] = v
    y = isht(c.to(DEV))[0].cpu().double()
    for i, (l, m, v) in enumerate(expect):
        err = _rel(y[i], torch.from_numpy(f[i]))
        assert err < TOL_T, (grid, l, m, v, err)
