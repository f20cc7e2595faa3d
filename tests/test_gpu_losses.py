"""GPU parity tests of makani_b200.GeometricH1Loss / degree_power (the spectral loss's transform).

1. against values and gradients recorded from the reference's own, unmodified GeometricH1Loss
   (makani/utils/losses.py:275-370; tests/golden/h1_loss_ref.npz, tools/make_golden.py) - every switch of the class,
   with and without the example mask;
2. the two kernels (sfno_sp_degree_power / _bwd) against the reference's arithmetic applied to the complex64 output of
   the product's own RealSHT - same transform on both sides, so this isolates the fused reduction and its adjoint;
3. against the fp64 CPU oracle on a grid that takes the tiled nlon = 480 FFT kernels;
4. at the reference's full size (721 x 1440, all 721 degrees and modes) through 2.

Tolerances: TOL = 2e-3 relative on loss values and 3e-3 relative L2 on gradients where the other side is fp32 / fp64
arithmetic (the Legendre contraction runs in TF32, as under the reference's allow_tf32); 5e-5 where both sides share
the product transform.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL, TOL_G, TOL_SAME = 2e-3, 3e-3, 5e-5
GOLD = os.path.join(os.path.dirname(__file__), "golden")

VARIANTS = {"rel": {}, "abs": {"absolute": True}, "rel_sq": {"squared": True},
            "abs_sq_mean": {"absolute": True, "squared": True, "size_average": True},
            "rel_mean": {"size_average": True}, "rel_none": {"reduction": False}, "rel_alpha": {"alpha": 0.2}}


@pytest.fixture(scope="module")
def mk(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import makani_b200
    return makani_b200


def _rel(a, b):
    a, b = a.detach().double(), b.detach().double()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def _formula(c):
    """losses.py:299-301 on a complex coefficient tensor"""
    p = torch.view_as_real(c)
    p = p[..., 0] ** 2 + p[..., 1] ** 2
    return p[..., :, 0] + 2 * torch.sum(p[..., :, 1:], dim=-1)


@pytest.mark.parametrize("tag,nlat,nlon", [("a", 13, 24), ("b", 25, 48)])
def test_h1_loss_matches_reference_class(mk, tag, nlat, nlon):
    z = np.load(os.path.join(GOLD, "h1_loss_ref.npz"))
    tar = torch.from_numpy(z[f"{tag}_tar"]).to(DEV)
    mask = torch.from_numpy(z[f"{tag}_mask"]).to(DEV)
    for name, kw in VARIANTS.items():
        mod = mk.GeometricH1Loss((nlat, nlon), **kw).to(DEV)
        assert mod.h1_weights.device.type == "cuda" and mod.sht.lmax == nlat
        for mtag, m in (("", None), ("_masked", mask)):
            prd = torch.from_numpy(z[f"{tag}_prd"]).to(DEV).requires_grad_(True)
            val = mod(prd, tar, m)
            val.sum().backward()
            want = torch.from_numpy(np.asarray(z[f"{tag}_{name}{mtag}"]))
            assert val.shape == want.shape, (name, mtag)
            np.testing.assert_allclose(val.detach().cpu().numpy(), want.numpy(), rtol=TOL, atol=1e-6,
                                       err_msg=name + mtag)
            assert _rel(prd.grad.cpu(), torch.from_numpy(z[f"{tag}_{name}{mtag}_dprd"])) < TOL_G, (name, mtag)


@pytest.mark.parametrize("nlat,nlon,lmax,mmax,B,Cn,dtype", [
    (31, 60, None, None, 2, 5, torch.float32),           # general DFT kernel, mmax = lmax = nlat
    (40, 96, 24, 33, 1, 7, torch.float32),               # mmax > lmax: columns above the diagonal carry no energy
    (120, 480, None, None, 2, 6, torch.float32),         # tiled nlon = 480 kernels (mmax 241 > lmax 120)
    (120, 480, 100, 64, 1, 9, torch.bfloat16),           # bf16 fields, odd channel count (pad channel in SP)
])
def test_degree_power_equals_the_formula_on_the_coefficients(mk, nlat, nlon, lmax, mmax, B, Cn, dtype):
    torch.manual_seed(1)
    sht = mk.RealSHT(nlat, nlon, lmax=lmax, mmax=mmax, grid="equiangular").to(DEV)
    x = torch.randn(B, Cn, nlat, nlon, device=DEV).to(dtype)
    g = torch.randn(B, Cn, sht.lmax, device=DEV)
    xa = x.clone().requires_grad_(True)
    ea = mk.losses.degree_power(xa, sht)
    (ea * g).sum().backward()
    xb = x.clone().requires_grad_(True)
    eb = _formula(sht(xb))
    (eb * g).sum().backward()
    assert ea.shape == (B, Cn, sht.lmax) and ea.dtype == torch.float32
    assert _rel(ea, eb) < TOL_SAME
    assert xa.grad.dtype == dtype
    assert _rel(xa.grad.float(), xb.grad.float()) < (TOL_SAME if dtype == torch.float32 else 1e-2)


def test_degree_power_without_grad_saves_nothing(mk):
    sht = mk.RealSHT(16, 32, grid="equiangular").to(DEV)
    x = torch.randn(1, 2, 16, 32, device=DEV)
    e = mk.losses.degree_power(x, sht)
    assert not e.requires_grad and e.shape == (1, 2, 16)
    with pytest.raises(RuntimeError):
        mk.losses.degree_power(x.cpu(), sht)
    with pytest.raises(TypeError):
        mk.losses.degree_power(x, torch.nn.Identity())


def test_h1_loss_against_the_oracle_on_the_fast_dft_grid(mk):
    from oracle import sht_oracle as O
    nlat, nlon, B, Cn = 61, 480, 2, 3
    torch.manual_seed(2)
    prd = torch.randn(B, Cn, nlat, nlon, dtype=torch.float64, requires_grad=True)
    tar = torch.randn(B, Cn, nlat, nlon, dtype=torch.float64)
    osht = O.RealSHT(nlat, nlon, grid="equiangular")
    want = O.geometric_h1_loss(prd, tar, osht)
    want.backward()
    mod = mk.GeometricH1Loss((nlat, nlon)).to(DEV)
    p = prd.detach().float().to(DEV).requires_grad_(True)
    got = mod(p, tar.float().to(DEV))
    got.backward()
    assert abs(float(got) - float(want)) < TOL * abs(float(want))
    assert _rel(p.grad.cpu(), prd.grad) < TOL_G


def test_degree_power_at_the_reference_resolution(mk):
    """721 x 1440 with all 721 degrees and longitude modes, as GeometricH1Loss builds its transform for the ERA5 grid
    (losses.py:291): the general DFT kernel at 202 KB of shared memory and a 1.5 GB Legendre table."""
    nlat, nlon, Cn = 721, 1440, 2
    torch.manual_seed(3)
    sht = mk.RealSHT(nlat, nlon, grid="equiangular").to(DEV)
    assert (sht.lmax, sht.mmax) == (721, 721)
    x = torch.randn(1, Cn, nlat, nlon, device=DEV)
    g = torch.rand(1, Cn, 721, device=DEV)
    xa = x.clone().requires_grad_(True)
    ea = mk.losses.degree_power(xa, sht)
    (ea * g).sum().backward()
    xb = x.clone().requires_grad_(True)
    eb = _formula(sht(xb))
    (eb * g).sum().backward()
    assert _rel(ea, eb) < TOL_SAME
    assert _rel(xa.grad, xb.grad) < TOL_SAME
    assert float(ea.min()) > 0
