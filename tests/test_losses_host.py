"""CPU test of the host side of makani_b200.GeometricH1Loss (makani/utils/losses.py:275-370): constructor, buffer, the
abs / rel / mask / reduction switches.  The CUDA reduction ``degree_power`` is replaced by the oracle's (fp32 CPU), so
what is compared with the values and gradients recorded from the reference's own class (tests/golden/h1_loss_ref.npz)
is the module logic around it; the kernels themselves are covered by tests/test_gpu_losses.py."""
import os

import numpy as np
import pytest
import torch

import makani_b200
from makani_b200 import losses
from oracle import sht_oracle as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")
VARIANTS = {"rel": {}, "abs": {"absolute": True}, "rel_sq": {"squared": True},
            "abs_sq_mean": {"absolute": True, "squared": True, "size_average": True},
            "rel_mean": {"size_average": True}, "rel_none": {"reduction": False}, "rel_alpha": {"alpha": 0.2}}


@pytest.mark.parametrize("tag,nlat,nlon", [("a", 13, 24), ("b", 25, 48)])
def test_module_logic_matches_reference_class(monkeypatch, tag, nlat, nlon):
    z = np.load(os.path.join(GOLD, "h1_loss_ref.npz"))
    osht = O.RealSHT(nlat, nlon, grid="equiangular", dtype=torch.float32)
    monkeypatch.setattr(losses, "degree_power", lambda x, sht: O.degree_power(x, osht))
    tar, mask = torch.from_numpy(z[f"{tag}_tar"]), torch.from_numpy(z[f"{tag}_mask"])
    for name, kw in VARIANTS.items():
        mod = makani_b200.GeometricH1Loss((nlat, nlon), **kw)
        assert (mod.sht.nlat, mod.sht.nlon, mod.sht.grid, mod.sht.lmax) == (nlat, nlon, "equiangular", osht.lmax)
        assert torch.equal(mod.h1_weights, torch.arange(nlat).float() * (torch.arange(nlat).float() + 1))
        for mtag, m in (("", None), ("_masked", mask)):
            prd = torch.from_numpy(z[f"{tag}_prd"]).requires_grad_(True)
            val = mod(prd, tar, m)
            val.sum().backward()
            np.testing.assert_allclose(val.detach().numpy(), z[f"{tag}_{name}{mtag}"], rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(prd.grad.numpy(), z[f"{tag}_{name}{mtag}_dprd"], rtol=1e-4, atol=1e-6)


def test_state_dict_and_cpu_refusal():
    mod = makani_b200.GeometricH1Loss((13, 24))
    assert list(mod.state_dict().keys()) == ["h1_weights"]          # the transform's tables are non-persistent
    with pytest.raises(RuntimeError, match="no CPU path"):
        mod(torch.randn(1, 1, 13, 24), torch.randn(1, 1, 13, 24))
    with pytest.raises(TypeError):
        losses.degree_power(torch.randn(1, 13, 24), torch.nn.Identity())
