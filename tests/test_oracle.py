"""CPU tests that pin the oracle (oracle/sht_oracle.py).

* contractions: against golden outputs of the reference's own contractions.py (tools/make_golden.py).
* quadrature:   against the reference's known-answer test (tests/test_metrics.py:37-45: the quadrature of
                the constant 1 over the sphere is 1 at rtol 1e-5 for (20,101)-style and (721,1440) grids)
                and numpy's Gauss-Legendre nodes.
* SHT values:   parity unpinned in the reference (torch-harmonics is not vendored); anchored on invariants:
                orthonormality vs scipy.special.sph_harm_y, exact round trip, projection, adjointness.
"""
import os

import numpy as np
import pytest
import torch

from oracle import sht_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name", ["a", "b"])
def test_contractions_match_reference_outputs(name):
    z = np.load(os.path.join(GOLD, f"contractions_{name}.npz"))
    x = torch.from_numpy(z["x"])
    y = O.contract_dhconv(x, torch.from_numpy(z["w_dhconv"]))
    assert torch.allclose(y, torch.from_numpy(z["y_dhconv"]), rtol=1e-5, atol=1e-5)
    y = O.contract_diagonal(x, torch.from_numpy(z["w_diagonal"]))
    assert torch.allclose(y, torch.from_numpy(z["y_diagonal"]), rtol=1e-5, atol=1e-5)
    assert torch.equal(O.contract(x, torch.from_numpy(z["w_dhconv"]), False, "dhconv"),
                       O.contract_dhconv(x, torch.from_numpy(z["w_dhconv"])))
    with pytest.raises(ValueError):
        O.contract(x, torch.from_numpy(z["w_dhconv"]), False, "nope")


@pytest.mark.parametrize("n", [20, 721])
def test_quadrature_known_answers(n):
    z = np.load(os.path.join(GOLD, "quadrature.npz"))
    x, w = O.legendre_gauss_weights(n, -1, 1)
    np.testing.assert_allclose(x, z[f"lg_nodes_{n}"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(w, z[f"lg_weights_{n}"], rtol=0, atol=1e-15)
    for fn in (O.legendre_gauss_weights, O.clenshaw_curtiss_weights, O.lobatto_weights):
        nodes, wts = fn(n, -1, 1)
        # reference GridQuadrature: q = w * 2*pi/nlon / (4*pi) summed over nlon points  ->  sum(w)/2 == 1
        np.testing.assert_allclose(wts.sum() / 2.0, 1.0, rtol=1e-5)
        assert np.all(np.diff(nodes) > 0) and nodes[0] >= -1 and nodes[-1] <= 1
        # exactness on low-degree polynomials: int x^2 = 2/3, int x^4 = 2/5
        np.testing.assert_allclose((wts * nodes ** 2).sum(), 2 / 3, rtol=1e-10)
        np.testing.assert_allclose((wts * nodes ** 4).sum(), 2 / 5, rtol=1e-10)


def test_split_shapes():
    assert O.compute_split_shapes(721, 4) == [181, 181, 181, 178]
    assert O.compute_split_shapes(240, 4) == [60] * 4
    assert O.compute_split_shapes(241, 2) == [121, 120]
    assert O.compute_split_shapes(384, 4) == [96] * 4
    assert O.compute_split_shapes(5, 4) == [1, 1, 1, 2]          # ceil-chunks would leave the last one empty
    assert O.compute_split_shapes(7, 1) == [7]


@pytest.mark.parametrize("grid,nlat", [("legendre-gauss", 24), ("equiangular", 49), ("lobatto", 25)])
def test_orthonormality_against_scipy(grid, nlat):
    # Clenshaw-Curtis with n nodes integrates degree <= n-1 exactly, so the equiangular case keeps lmax at n/2
    from scipy.special import sph_harm_y
    nlon = 48
    sht = O.RealSHT(nlat, nlon, lmax=24, mmax=25, grid=grid)
    cost, _, _ = O.grid_nodes_weights(nlat, grid)
    theta = np.flip(np.arccos(cost))
    phi = np.linspace(0, 2 * np.pi, nlon, endpoint=False)
    lmax = sht.lmax
    for (l, m) in [(0, 0), (1, 0), (1, 1), (7, 3), (lmax // 2, 1), (lmax // 2, lmax // 2), (10, 10)]:
        Y = sph_harm_y(l, m, theta[:, None], phi[None, :])
        field = Y.real if m == 0 else 2.0 * Y.real          # 2 Re Y_l^m = Y_l^m + conj -> coefficient 1 at (l, m)
        c = sht(torch.from_numpy(np.ascontiguousarray(field)))
        expect = torch.zeros_like(c)
        expect[l, m] = 1.0
        assert (c - expect).abs().max() < 1e-10, (grid, l, m, float((c - expect).abs().max()))


@pytest.mark.parametrize("nlat,nlon,grid,L,M", [(24, 48, "legendre-gauss", 24, 25), (49, 72, "equiangular", 24, 25),
                                                (120, 240, "legendre-gauss", 120, 121)])
def test_roundtrip_projection_adjoint(nlat, nlon, grid, L, M):
    torch.manual_seed(0)
    sht = O.RealSHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
    isht = O.InverseRealSHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
    c = torch.randn(3, L, M, dtype=torch.complex128) * torch.tril(torch.ones(L, M))
    c[..., 0] = c[..., 0].real.to(torch.complex128)
    if nlon % 2 == 0 and M == nlon // 2 + 1:
        c[..., M - 1] = c[..., M - 1].real.to(torch.complex128)
    x = isht(c)
    assert (sht(x) - c).abs().max() < 1e-11                    # sht o isht = id on band-limited input
    f = torch.randn(3, nlat, nlon, dtype=torch.float64)
    p = isht(sht(f))
    assert (isht(sht(p)) - p).abs().max() < 1e-11              # isht o sht is a projection
    # fp32 version of the oracle stays within 1e-5 of fp64
    sht32 = O.RealSHT(nlat, nlon, lmax=L, mmax=M, grid=grid, dtype=torch.float32)
    assert O.rel_l2(sht32(f.float()), sht(f)) < 1e-5


def test_golden_regression_small():
    z = np.load(os.path.join(GOLD, "sht_small.npz"))
    for tag, (nlat, nlon, grid) in {"lg": (12, 24, "legendre-gauss"), "eq": (13, 24, "equiangular"),
                                    "lob": (9, 16, "lobatto")}.items():
        sht = O.RealSHT(nlat, nlon, grid=grid)
        isht = O.InverseRealSHT(nlat, nlon, grid=grid)
        c = sht(torch.from_numpy(z[f"{tag}_x"]))
        np.testing.assert_allclose(c.numpy(), z[f"{tag}_c"], rtol=0, atol=1e-13)
        np.testing.assert_allclose(isht(c).numpy(), z[f"{tag}_xr"], rtol=0, atol=1e-13)


def test_spectral_conv_oracle_matches_composition():
    torch.manual_seed(1)
    fwd = O.RealSHT(16, 32, grid="legendre-gauss")
    inv = O.InverseRealSHT(16, 32, grid="legendre-gauss")
    w = O.init_spectral_weight(4, 6, 16, operator_type="dhconv").to(torch.complex128)
    x = torch.randn(2, 4, 16, 32, dtype=torch.float64)
    y, res = O.spectral_conv_forward(x, fwd, inv, w, "dhconv", compute_dtype=torch.float64)
    assert res is x
    y2 = inv(torch.einsum("bixy,iox->boxy", fwd(x), w))
    assert torch.allclose(y, y2)
    inv2 = O.InverseRealSHT(24, 48, lmax=16, mmax=17, grid="equiangular")
    y3, res3 = O.spectral_conv_forward(x, fwd, inv2, w, "dhconv", compute_dtype=torch.float64)
    assert res3.shape == (2, 4, 24, 48) and y3.shape == (2, 6, 24, 48)
    # weight init statistics of spectral_convolution.py:97-102
    w0 = O.init_spectral_weight(64, 64, 32, operator_type="dhconv")
    assert abs(float(w0[..., 1:].abs().pow(2).mean()) - 1 / 64) < 2e-3
    assert abs(float(w0[..., 0].abs().pow(2).mean()) - 2 / 64) < 6e-3


@pytest.mark.parametrize("name", ["internal", "down", "up", "diag"])
def test_spectral_conv_matches_reference_module(name):
    """oracle.spectral_conv_forward against outputs AND gradients of the reference's own, unmodified SpectralConv
    (spectral_convolution.py:43-148; fixtures made by tools/make_golden.py through tests/refharness.py).  Both sides
    use the oracle's SHT arithmetic in fp32, so this pins the orchestration: fp32 forcing, residual re-projection,
    contraction dispatch, bias, result dtype."""
    from golden_cases import SPECTRAL_CONV_CASES, load
    gin, gout, L, M, B, Ci, Co, op, bias = SPECTRAL_CONV_CASES[name]
    z = load(name)
    fwd = O.RealSHT(gin[0], gin[1], lmax=L, mmax=M, grid=gin[2], dtype=torch.float32)
    inv = O.InverseRealSHT(gout[0], gout[1], lmax=L, mmax=M, grid=gout[2], dtype=torch.float32)
    x = torch.from_numpy(z["x"]).requires_grad_(True)
    w = torch.from_numpy(z["weight"]).requires_grad_(True)
    b = torch.from_numpy(z["bias"]).requires_grad_(True) if bias else None
    assert x.shape == (B, Ci, gin[0], gin[1]) and w.shape[:2] == (Ci, Co)
    y, res = O.spectral_conv_forward(x, fwd, inv, w, op, bias=b, compute_dtype=torch.float32)
    assert y.dtype == torch.float32 and tuple(y.shape) == z["y"].shape
    assert O.rel_l2(y.detach(), torch.from_numpy(z["y"])) < 1e-6
    scale_residual = bool(z["scale_residual"])
    assert scale_residual == (gin != gout)
    loss = (y * torch.from_numpy(z["gy"])).sum()
    if scale_residual:
        assert O.rel_l2(res.detach(), torch.from_numpy(z["residual"])) < 1e-6
        loss = loss + (res * torch.from_numpy(z["gr"])).sum()
    else:
        assert res is x
    loss.backward()
    assert O.rel_l2(x.grad, torch.from_numpy(z["dx"])) < 1e-6
    dw = torch.from_numpy(z["dweight"])
    assert float((w.grad - dw).norm() / dw.norm()) < 1e-6
    if bias:
        db = torch.from_numpy(z["dbias"])
        assert float((b.grad - db).norm() / db.norm()) < 1e-6


H1_VARIANTS = {"rel": {}, "abs": {"absolute": True}, "rel_sq": {"squared": True},
               "abs_sq_mean": {"absolute": True, "squared": True, "size_average": True},
               "rel_mean": {"size_average": True}, "rel_none": {"reduction": False}, "rel_alpha": {"alpha": 0.2}}


@pytest.mark.parametrize("tag,nlat,nlon", [("a", 13, 24), ("b", 25, 48)])
def test_h1_loss_matches_reference_class(tag, nlat, nlon):
    """oracle.geometric_h1_loss against values AND gradients recorded from the reference's own GeometricH1Loss
    (makani/utils/losses.py:275-370, tools/make_golden.py), every switch, with and without the example mask."""
    z = np.load(os.path.join(GOLD, "h1_loss_ref.npz"))
    sht = O.RealSHT(nlat, nlon, grid="equiangular", dtype=torch.float32)
    tar, mask = torch.from_numpy(z[f"{tag}_tar"]), torch.from_numpy(z[f"{tag}_mask"])
    for name, kw in H1_VARIANTS.items():
        for mtag, m in (("", None), ("_masked", mask)):
            prd = torch.from_numpy(z[f"{tag}_prd"]).requires_grad_(True)
            val = O.geometric_h1_loss(prd, tar, sht, mask=m, **kw)
            val.sum().backward()
            np.testing.assert_allclose(val.detach().numpy(), z[f"{tag}_{name}{mtag}"], rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(prd.grad.numpy(), z[f"{tag}_{name}{mtag}_dprd"], rtol=1e-4, atol=1e-6)


def test_degree_power_is_the_quadrature_of_the_square():
    """Parseval on the sphere: for a band-limited field the sum over l of the per-degree energy equals the integral of
    x^2 (Gauss-Legendre quadrature is exact for it) - anchors the 'once at m = 0, twice elsewhere' weighting."""
    nlat, nlon, L = 24, 48, 12
    sht = O.RealSHT(nlat, nlon, lmax=L, mmax=L, grid="legendre-gauss")
    isht = O.InverseRealSHT(nlat, nlon, lmax=L, mmax=L, grid="legendre-gauss")
    g = torch.Generator().manual_seed(5)
    c = torch.complex(torch.randn(2, L, L, dtype=torch.float64, generator=g),
                      torch.randn(2, L, L, dtype=torch.float64, generator=g))
    c = c * (torch.arange(L)[None, :] <= torch.arange(L)[:, None])            # m <= l
    c[..., 0] = c[..., 0].real.to(torch.complex128)
    x = isht(c)
    _, w, _ = O.grid_nodes_weights(nlat, "legendre-gauss")
    integral = (x ** 2 * torch.from_numpy(np.ascontiguousarray(w))[:, None]).sum(dim=(-2, -1)) * 2 * np.pi / nlon
    np.testing.assert_allclose(O.degree_power(x, sht).sum(-1).numpy(), integral.numpy(), rtol=1e-10)
