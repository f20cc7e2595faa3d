"""Shared description of the reference-generated SpectralConv fixtures (tools/make_golden.py: outputs and gradients of
the reference's own, unmodified SpectralConv with the CPU oracle standing in for torch-harmonics)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# name: (in grid, out grid, lmax, mmax, B, Cin, Cout, operator, bias)   - keep in sync with tools/make_golden.py
SPECTRAL_CONV_CASES = {
    "internal": ((16, 32, "legendre-gauss"), (16, 32, "legendre-gauss"), 16, 17, 2, 8, 8, "dhconv", False),
    "down": ((33, 64, "equiangular"), (16, 32, "legendre-gauss"), 16, 17, 1, 8, 12, "dhconv", False),
    "up": ((16, 32, "legendre-gauss"), (33, 64, "equiangular"), 16, 17, 2, 6, 8, "dhconv", "constant"),
    "diag": ((12, 24, "legendre-gauss"), (12, 24, "legendre-gauss"), 11, 11, 2, 4, 6, "diagonal", "position"),
}


def load(name):
    return np.load(os.path.join(GOLDEN, f"spectral_conv_ref_{name}.npz"))
