"""TEST INFRASTRUCTURE: load UNMODIFIED reference modules from /root/reference in the build container.

`import makani` itself fails here (its __init__ pulls the trainer; ruamel, h5py, wandb-era dependencies, modulus,
tensorly, tltorch, torch-harmonics and parameterized are not installed).  The spectral path's own files need far less,
so this harness registers the minimum in ``sys.modules`` and lets Python import the reference files where they lie:

* ``makani``, ``makani.models``, ``makani.utils``  path-only package stubs (no __init__ executed); every other
  ``makani.*`` module is the reference's own file, unmodified
* ``makani.utils.comm``                  -> ``makani_b200.comm`` (the product's registry)
* ``torch_harmonics``                    -> the product shim (``th="product"``) or nn.Module wrappers around the CPU
                                            oracle (``th="oracle"``; the distributed sub-module is the product's in
                                            both cases - its transposes are device agnostic)
* ``modulus.distributed.{utils,mappings}`` -> ``makani_b200.distributed`` / ``.mappings``; ``modulus.Module`` and
  ``modulus.models.meta.ModelMetaData`` inert stand-ins (sfnonet.py:642-672 only registers the model with them)
* ``tensorly`` / ``tltorch``             import-time stand-ins (only the factorized variants, out of scope, use them)
* ``parameterized``                      ``parameterized.expand`` for the reference's unittest files

Nothing here is used by the product, by bench.py or on the GPU box (where /root/reference does not exist):
``available()`` is False there and the tests that need the harness skip.
"""
from __future__ import annotations

import dataclasses
import functools
import importlib
import os
import sys
import types

import torch
import torch.nn as nn

REF_ROOT = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "makani", "models", "networks", "sfnonet.py"))


def _module(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    parent, _, leaf = name.rpartition(".")
    if parent and parent in sys.modules:
        setattr(sys.modules[parent], leaf, m)
    return m


def _path_package(name: str, path: str) -> types.ModuleType:
    m = _module(name)
    m.__path__ = [path]
    m.__package__ = name
    return m


# ------------------------------------------------------------------------------------------ torch_harmonics (oracle)
def _oracle_th():
    from oracle import sht_oracle as O
    import makani_b200.distributed as thd
    from makani_b200 import legendre, quadrature

    class RealSHT(nn.Module):
        def __init__(self, nlat, nlon, lmax=None, mmax=None, grid="lobatto", norm="ortho", csphase=True):
            super().__init__()
            self._o = O.RealSHT(nlat, nlon, lmax=lmax, mmax=mmax, grid=grid, norm=norm, csphase=csphase)
            self.nlat, self.nlon, self.grid, self.norm, self.csphase = nlat, nlon, grid, norm, csphase
            self.lmax, self.mmax = self._o.lmax, self._o.mmax
            self.register_buffer("weights", self._o.weights.clone(), persistent=False)

        def forward(self, x):
            self._o.weights = self.weights
            return self._o(x)

    class InverseRealSHT(nn.Module):
        def __init__(self, nlat, nlon, lmax=None, mmax=None, grid="lobatto", norm="ortho", csphase=True):
            super().__init__()
            self._o = O.InverseRealSHT(nlat, nlon, lmax=lmax, mmax=mmax, grid=grid, norm=norm, csphase=csphase)
            self.nlat, self.nlon, self.grid, self.norm, self.csphase = nlat, nlon, grid, norm, csphase
            self.lmax, self.mmax = self._o.lmax, self._o.mmax
            self.register_buffer("pct", self._o.pct.clone(), persistent=False)

        def forward(self, x):
            self._o.pct = self.pct
            return self._o(x)

    th = _module("torch_harmonics", RealSHT=RealSHT, InverseRealSHT=InverseRealSHT, __version__="0.6.5+oracle")
    th.__path__ = []
    th.distributed = thd
    th.quadrature = quadrature
    th.legendre = legendre
    sys.modules["torch_harmonics.distributed"] = thd
    sys.modules["torch_harmonics.quadrature"] = quadrature
    sys.modules["torch_harmonics.legendre"] = legendre
    return th


# ------------------------------------------------------------------------------------------ third-party stand-ins
def _install_third_party():
    import makani_b200.distributed as thd
    from makani_b200.distributed import mappings

    tl = _module("tensorly", set_backend=lambda name: None, ndim=lambda x: x.dim(), einsum=torch.einsum)
    tl.__path__ = []

    class FactorizedTensor:                                   # tltorch.factorized_tensors.core.FactorizedTensor
        @classmethod
        def new(cls, *a, **k):
            raise NotImplementedError("tltorch is not installed: factorized spectral weights are out of scope")

    _module("tltorch").__path__ = []
    _module("tltorch.factorized_tensors").__path__ = []
    _module("tltorch.factorized_tensors.core", FactorizedTensor=FactorizedTensor)

    @dataclasses.dataclass
    class ModelMetaData:
        name: str = "model"
        jit: bool = False
        cuda_graphs: bool = False
        amp_cpu: bool = False
        amp_gpu: bool = False

    class Module(nn.Module):
        @classmethod
        def from_torch(cls, torch_model_class, meta=None):
            return torch_model_class

    mod = _module("modulus", Module=Module)
    mod.__path__ = []
    _module("modulus.models").__path__ = []
    _module("modulus.models.meta", ModelMetaData=ModelMetaData)
    _module("modulus.distributed").__path__ = []
    _module("modulus.distributed.utils", compute_split_shapes=thd.compute_split_shapes,
            split_tensor_along_dim=thd.split_tensor_along_dim)
    _module("modulus.distributed.mappings", **{n: getattr(mappings, n) for n in mappings.__all__})

    class parameterized:                                      # noqa: N801  (the package exposes this lower-case class)
        @staticmethod
        def expand(cases):
            def deco(fn):
                frame_locals = sys._getframe(1).f_locals       # the TestCase class body being executed
                for i, case in enumerate(cases):
                    args = tuple(case)

                    @functools.wraps(fn)
                    def one(self, _args=args):
                        return fn(self, *_args)
                    one.__name__ = f"{fn.__name__}_{i}"
                    frame_locals[one.__name__] = one
                return None                                    # the undecorated name is removed, like the package does
            return deco

    _module("parameterized", parameterized=parameterized)


_INSTALLED = None


def install(th: str = "oracle"):
    """Register everything; returns the ``torch_harmonics`` module in use.  Idempotent per process for one ``th``."""
    global _INSTALLED
    if not available():
        raise RuntimeError("refharness: /root/reference is not present (build container only)")
    if _INSTALLED is not None:
        if _INSTALLED[0] != th:
            raise RuntimeError(f"refharness already installed with th={_INSTALLED[0]!r}")
        return _INSTALLED[1]
    import makani_b200
    from makani_b200 import comm

    _install_third_party()
    if th == "oracle":
        thm = _oracle_th()
    elif th == "product":
        thm = makani_b200.install_torch_harmonics_shim(force=True)
    else:
        raise ValueError(th)
    pkg = os.path.join(REF_ROOT, "makani")
    _path_package("makani", pkg)
    _path_package("makani.models", os.path.join(pkg, "models"))          # its __init__ imports the model registry
    _path_package("makani.utils", os.path.join(pkg, "utils"))
    sys.modules["makani.utils.comm"] = comm
    sys.modules["makani.utils"].comm = comm
    _INSTALLED = (th, thm)
    return thm


def load(name: str):
    """``importlib.import_module`` of a reference module, e.g. ``makani.models.common.spectral_convolution``."""
    assert _INSTALLED is not None, "call refharness.install() first"
    return importlib.import_module(name)


def load_file(path: str, name: str):
    """Execute a reference file that is not part of the package (its tests/), by path relative to the reference."""
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF_ROOT, path))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod
