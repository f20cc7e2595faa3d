"""Quadrature nodes / weights on [-1, 1] for the latitude grids of the SHT (host side, fp64 numpy).

Mirrors the ``torch_harmonics.quadrature`` API that the reference consumes at
makani/utils/grids.py:32,77,83 and makani/utils/losses.py:31: every function returns
``(nodes_cos_theta_ascending, weights)`` as numpy arrays.
"""
from __future__ import annotations

import numpy as np

__all__ = ["legendre_gauss_weights", "clenshaw_curtiss_weights", "lobatto_weights", "fejer2_weights",
           "trapezoidal_weights", "precompute_latitudes", "precompute_longitudes"]


def _affine(nodes, weights, a, b):
    half = 0.5 * (b - a)
    return half * nodes + 0.5 * (b + a), half * weights


def legendre_gauss_weights(n, a=-1.0, b=1.0):
    """Gauss-Legendre rule: exact for polynomials of degree 2n-1."""
    nodes, weights = np.polynomial.legendre.leggauss(n)
    return _affine(nodes, weights, a, b)


def clenshaw_curtiss_weights(n, a=-1.0, b=1.0):
    """Clenshaw-Curtis rule on the n Chebyshev extrema (poles included), via one length-(n-1) inverse FFT
    (Waldvogel's construction)."""
    if n < 2:
        raise ValueError("clenshaw_curtiss_weights needs n >= 2")
    nodes = np.cos(np.linspace(np.pi, 0.0, n))
    if n == 2:
        weights = np.ones(2)
    else:
        p = n - 1
        odd = np.arange(1, p, 2)
        n_odd = odd.size
        n_rest = p - n_odd
        v = np.concatenate([2.0 / odd / (odd - 2), 1.0 / odd[-1:], np.zeros(n_rest)])
        v = -v[:-1] - v[-1:0:-1]
        g = -np.ones(p)
        g[n_odd] += p
        g[n_rest] += p
        g /= p * p - 1 + (p % 2)
        weights = np.fft.ifft(v + g).real
        weights = np.concatenate([weights, weights[:1]])
    return _affine(nodes, weights, a, b)


def fejer2_weights(n, a=-1.0, b=1.0):
    """Fejer's second rule (Chebyshev extrema, poles carry zero weight)."""
    if n < 3:
        raise ValueError("fejer2_weights needs n >= 3")
    nodes = np.cos(np.linspace(np.pi, 0.0, n))
    p = n - 1
    odd = np.arange(1, p, 2)
    n_odd = odd.size
    n_rest = p - n_odd
    v = np.concatenate([2.0 / odd / (odd - 2), 1.0 / odd[-1:], np.zeros(n_rest)])
    v = -v[:-1] - v[-1:0:-1]
    weights = np.fft.ifft(v).real
    weights = np.concatenate([weights, weights[:1]])
    return _affine(nodes, weights, a, b)


def lobatto_weights(n, a=-1.0, b=1.0, tol=1e-16, maxiter=100):
    """Gauss-Lobatto rule: Newton iteration on the extrema of P_{n-1}, started from Chebyshev extrema."""
    x = -np.cos(np.pi * np.arange(n) / (n - 1))
    leg = np.zeros((n, n))
    for _ in range(maxiter):
        x_prev = x
        leg[:, 0] = 1.0
        leg[:, 1] = x
        for k in range(2, n):
            leg[:, k] = ((2 * k - 1) * x * leg[:, k - 1] - (k - 1) * leg[:, k - 2]) / k
        x = x_prev - (x * leg[:, n - 1] - leg[:, n - 2]) / (n * leg[:, n - 1])
        if np.max(np.abs(x - x_prev)) < tol:
            break
    weights = 2.0 / ((n * (n - 1)) * leg[:, n - 1] ** 2)
    return _affine(x, weights, a, b)


def trapezoidal_weights(n, a=-1.0, b=1.0, periodic=False):
    nodes = np.linspace(a, b, n, endpoint=periodic)
    h = (b - a) / n if periodic else (b - a) / (n - 1)
    weights = np.full(n, h)
    if not periodic:
        weights[0] *= 0.5
        weights[-1] *= 0.5
    return nodes, weights


def precompute_latitudes(nlat, grid="equiangular"):
    """Colatitudes (north -> south) and quadrature weights for a named latitude grid."""
    if grid == "legendre-gauss":
        cost, w = legendre_gauss_weights(nlat)
    elif grid == "lobatto":
        cost, w = lobatto_weights(nlat)
    elif grid == "equiangular":
        cost, w = clenshaw_curtiss_weights(nlat)
    else:
        raise ValueError(f"Unknown grid {grid}")
    return np.flip(np.arccos(cost)).copy(), np.flip(w).copy()


def precompute_longitudes(nlon):
    return np.linspace(0.0, 2.0 * np.pi, nlon, endpoint=False)
