"""Generate the golden fixtures under tests/golden/ .  Run in the build container only (needs /root/reference):

    python tools/make_golden.py

1. contractions_*.npz  - outputs of the REFERENCE's own makani/models/common/contractions.py (loaded by file
   path; it imports only torch) on seeded inputs: pins oracle.contract_dhconv / contract_diagonal.
2. quadrature.npz      - the reference's known-answer for quadrature weights (tests/test_metrics.py:37-45:
   integral of 1 over the sphere == 1 after the GridQuadrature normalisation of makani/utils/grids.py:63-97,
   restated here because grids.py imports torch_harmonics which is absent) and numpy's Gauss-Legendre rule.
4. spectral_conv_ref_*.npz - outputs and gradients of the REFERENCE's own, unmodified ``SpectralConv``
   (makani/models/common/spectral_convolution.py:43-148, loaded through tests/refharness.py) on seeded inputs, with the
   CPU oracle standing in for torch-harmonics (fp32, as ``sfnonet.py:536-539`` builds the transforms with ``.float()``):
   pins the module's orchestration - fp32 forcing, residual re-projection, contraction dispatch, bias, dtype of the
   results - for oracle.spectral_conv_forward (CPU) and makani_b200.SpectralConv (GPU).
5. h1_loss_ref.npz     - values and gradients of the REFERENCE's own, unmodified ``GeometricH1Loss``
   (makani/utils/losses.py:275-370, loaded through tests/refharness.py) with the CPU oracle standing in for
   torch-harmonics, for every combination of its switches: pins oracle.geometric_h1_loss (CPU) and
   makani_b200.GeometricH1Loss (GPU).  ``python tools/make_golden.py h1`` writes this file only.
3. sht_small.npz       - oracle outputs on small grids (NOT reference outputs: torch-harmonics is not available;
   recorded so a later change of the oracle or of the kernels is caught as a regression).
"""
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference/makani/models/common/contractions.py"
OUT = os.path.join(ROOT, "tests", "golden")


def load_ref_contractions():
    spec = importlib.util.spec_from_file_location("ref_contractions", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


SPECTRAL_CONV_CASES = {
    # name: (in grid, out grid, lmax, mmax, B, Cin, Cout, operator, bias)
    "internal": ((16, 32, "legendre-gauss"), (16, 32, "legendre-gauss"), 16, 17, 2, 8, 8, "dhconv", False),
    "down": ((33, 64, "equiangular"), (16, 32, "legendre-gauss"), 16, 17, 1, 8, 12, "dhconv", False),
    "up": ((16, 32, "legendre-gauss"), (33, 64, "equiangular"), 16, 17, 2, 6, 8, "dhconv", "constant"),
    # diagonal: the reference's init broadcasts its per-l scale over the LAST (m) axis, so it only constructs with L == M
    "diag": ((12, 24, "legendre-gauss"), (12, 24, "legendre-gauss"), 11, 11, 2, 4, 6, "diagonal", "position"),
}


def make_spectral_conv_reference():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refharness as H
    th = H.install("oracle")
    sc = H.load("makani.models.common.spectral_convolution")
    assert sc.__file__.startswith("/root/reference/")
    for name, (gin, gout, L, M, B, Ci, Co, op, bias) in SPECTRAL_CONV_CASES.items():
        torch.manual_seed(sum(map(ord, name)))
        fwd = th.RealSHT(gin[0], gin[1], lmax=L, mmax=M, grid=gin[2]).float()
        inv = th.InverseRealSHT(gout[0], gout[1], lmax=L, mmax=M, grid=gout[2]).float()
        mod = sc.SpectralConv(fwd, inv, Ci, Co, operator_type=op, bias=bias)
        if bias:
            with torch.no_grad():
                mod.bias.copy_(0.3 * torch.randn_like(mod.bias))
        x = torch.randn(B, Ci, gin[0], gin[1], requires_grad=True)
        y, res = mod(x)
        gy = torch.randn_like(y)
        gr = torch.randn_like(res)
        loss = (y * gy).sum() + ((res * gr).sum() if mod.scale_residual else 0.0)
        loss.backward()
        out = dict(x=x.detach().numpy(), weight=mod.weight.detach().numpy(), y=y.detach().numpy(),
                   residual=res.detach().numpy(), gy=gy.numpy(), gr=gr.numpy(), dx=x.grad.numpy(),
                   dweight=mod.weight.grad.numpy(), scale_residual=np.array(mod.scale_residual))
        if bias:
            out["bias"] = mod.bias.detach().numpy()
            out["dbias"] = mod.bias.grad.numpy()
        np.savez_compressed(os.path.join(OUT, f"spectral_conv_ref_{name}.npz"), **out)


H1_GRIDS = {"a": (13, 24, 2, 3), "b": (25, 48, 3, 4)}          # nlat, nlon, B, C
H1_VARIANTS = {"rel": {}, "abs": {"absolute": True}, "rel_sq": {"squared": True},
               "abs_sq_mean": {"absolute": True, "squared": True, "size_average": True},
               "rel_mean": {"size_average": True}, "rel_none": {"reduction": False}, "rel_alpha": {"alpha": 0.2}}


def make_h1_loss_reference():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refharness as H
    H.install("oracle")
    losses = H.load("makani.utils.losses")
    assert losses.__file__.startswith("/root/reference/")
    out = {}
    for tag, (nlat, nlon, B, Cn) in H1_GRIDS.items():
        torch.manual_seed(7 + nlat)
        prd = torch.randn(B, Cn, nlat, nlon, requires_grad=True)
        tar = torch.randn(B, Cn, nlat, nlon)
        mask = (torch.arange(B) % 2 == 0).float()
        out[f"{tag}_prd"], out[f"{tag}_tar"], out[f"{tag}_mask"] = prd.detach().numpy(), tar.numpy(), mask.numpy()
        for name, kw in H1_VARIANTS.items():
            mod = losses.GeometricH1Loss((nlat, nlon), **kw)
            for mtag, m in (("", None), ("_masked", mask)):
                prd.grad = None
                val = mod(prd, tar, m)
                val.sum().backward()
                out[f"{tag}_{name}{mtag}"] = val.detach().numpy()
                out[f"{tag}_{name}{mtag}_dprd"] = prd.grad.numpy().copy()
    np.savez_compressed(os.path.join(OUT, "h1_loss_ref.npz"), **out)


def main():
    os.makedirs(OUT, exist_ok=True)
    if sys.argv[1:] == ["h1"]:
        make_h1_loss_reference()
        print("wrote h1_loss_ref.npz")
        return
    ref = load_ref_contractions()
    g = torch.Generator().manual_seed(1234)
    for name, (B, Ci, Co, L, M) in {"a": (2, 6, 10, 12, 13), "b": (1, 16, 16, 20, 21)}.items():
        x = torch.randn(B, Ci, L, M, dtype=torch.complex64, generator=g)
        w_dh = torch.randn(Ci, Co, L, dtype=torch.complex64, generator=g)
        w_dg = torch.randn(Ci, Co, L, M, dtype=torch.complex64, generator=g)
        y_dh = ref._contract_dhconv(x, w_dh)
        y_dg = ref._contract_diagonal(x, w_dg)
        np.savez_compressed(os.path.join(OUT, f"contractions_{name}.npz"), x=x.numpy(), w_dhconv=w_dh.numpy(),
                            w_diagonal=w_dg.numpy(), y_dhconv=y_dh.numpy(), y_diagonal=y_dg.numpy())

    from oracle import sht_oracle as O
    quad = {}
    for n in (20, 721):
        x, w = np.polynomial.legendre.leggauss(n)
        quad[f"lg_nodes_{n}"], quad[f"lg_weights_{n}"] = x, w
    np.savez_compressed(os.path.join(OUT, "quadrature.npz"), **quad)

    small = {}
    g = torch.Generator().manual_seed(99)
    for tag, (nlat, nlon, grid) in {"lg": (12, 24, "legendre-gauss"), "eq": (13, 24, "equiangular"),
                                    "lob": (9, 16, "lobatto")}.items():
        sht = O.RealSHT(nlat, nlon, grid=grid)
        isht = O.InverseRealSHT(nlat, nlon, grid=grid)
        x = torch.randn(2, nlat, nlon, dtype=torch.float64, generator=g)
        c = sht(x)
        small[f"{tag}_x"] = x.numpy()
        small[f"{tag}_c"] = c.numpy()
        small[f"{tag}_xr"] = isht(c).numpy()
    np.savez_compressed(os.path.join(OUT, "sht_small.npz"), **small)
    make_spectral_conv_reference()
    make_h1_loss_reference()
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
