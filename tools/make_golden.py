"""Generate the golden fixtures under tests/golden/ .  Run in the build container only (needs /root/reference):

    python tools/make_golden.py

1. contractions_*.npz  - outputs of the REFERENCE's own makani/models/common/contractions.py (loaded by file
   path; it imports only torch) on seeded inputs: pins oracle.contract_dhconv / contract_diagonal.
2. quadrature.npz      - the reference's known-answer for quadrature weights (tests/test_metrics.py:37-45:
   integral of 1 over the sphere == 1 after the GridQuadrature normalisation of makani/utils/grids.py:63-97,
   restated here because grids.py imports torch_harmonics which is absent) and numpy's Gauss-Legendre rule.
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


def main():
    os.makedirs(OUT, exist_ok=True)
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
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
