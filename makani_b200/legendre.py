"""Host-side fp64 precompute of the orthonormal associated-Legendre table used by the SHT kernels.

``legpoly(mmax, lmax, theta)[m, l, k] = Pbar_l^m(cos theta_k)`` with the Condon-Shortley phase - the same
table torch-harmonics' ``_precompute_legpoly`` yields (the reference builds its transforms from it,
makani/models/networks/sfnonet.py:536-539).  Computed once per transform on the host, then uploaded to
the device by ``sfno_plan_create``.  The recurrence runs over degree l with all orders m vectorised.
"""
from __future__ import annotations

import numpy as np

__all__ = ["legpoly", "precompute_legpoly"]


def legpoly(mmax, lmax, theta, norm="ortho", inverse=False, csphase=True):
    theta = np.asarray(theta, dtype=np.float64)
    n = max(mmax, lmax)
    x = np.cos(theta)
    tab = np.zeros((n, n, theta.size), dtype=np.float64)      # [m, l, k]

    scale = 1.0 if norm == "ortho" else np.sqrt(4.0 * np.pi)
    if inverse:
        scale = 1.0 / scale
    tab[0, 0] = scale / np.sqrt(4.0 * np.pi)

    # sectoral (m == l) and first off-sectoral (m == l-1) terms
    for l in range(1, n):
        tab[l - 1, l] = np.sqrt(2 * l + 1) * x * tab[l - 1, l - 1]
        tab[l, l] = np.sqrt((2 * l + 1) * (1 + x) * (1 - x) / 2 / l) * tab[l - 1, l - 1]

    # three-term recurrence in l, all m <= l-2 at once
    for l in range(2, n):
        m = np.arange(0, l - 1, dtype=np.float64)
        a = np.sqrt((2 * l - 1) / (l - m) * (2 * l + 1) / (l + m))
        b = np.sqrt((l + m - 1) / (l - m) * (2 * l + 1) / (2 * l - 3) * (l - m - 1) / (l + m))
        tab[: l - 1, l] = x[None, :] * a[:, None] * tab[: l - 1, l - 1] - b[:, None] * tab[: l - 1, l - 2]

    if norm == "schmidt":
        deg = np.sqrt(2.0 * np.arange(n) + 1.0)[None, :, None]
        tab = tab * deg if inverse else tab / deg

    tab = tab[:mmax, :lmax]
    if csphase:
        tab[1::2] *= -1.0
    return np.ascontiguousarray(tab)


# torch-harmonics spelling
precompute_legpoly = legpoly
