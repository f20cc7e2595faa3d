"""Import shim: lets unmodified reference code (`import torch_harmonics as th`,
`import torch_harmonics.distributed as thd`, `from torch_harmonics.quadrature import ...`) resolve to this
package.  See INTEGRATION.md.  Call sites: makani/models/networks/sfnonet.py:32-33,
makani/models/common/spectral_convolution.py:39-40, makani/utils/grids.py:19, makani/utils/losses.py:30-31,
makani/mpu/layers.py:34-35.
"""
from __future__ import annotations

import sys
import types


def install_torch_harmonics_shim(force: bool = False):
    """Register `torch_harmonics`, `torch_harmonics.distributed`, `.quadrature`, `.legendre` in sys.modules."""
    if "torch_harmonics" in sys.modules and not force:
        return sys.modules["torch_harmonics"]
    from . import distributed, legendre, quadrature, sht

    th = types.ModuleType("torch_harmonics")
    th.__doc__ = "makani_b200 stand-in for torch_harmonics (SFNO spectral hot path only)"
    th.RealSHT = sht.RealSHT
    th.InverseRealSHT = sht.InverseRealSHT
    th.quadrature = quadrature
    th.legendre = legendre
    th.distributed = distributed
    th.__version__ = "0.6.5+makani_b200"
    th.__path__ = []          # mark as package so submodule imports resolve through sys.modules
    sys.modules["torch_harmonics"] = th
    sys.modules["torch_harmonics.quadrature"] = quadrature
    sys.modules["torch_harmonics.legendre"] = legendre
    sys.modules["torch_harmonics.distributed"] = distributed
    sys.modules["torch_harmonics.sht"] = sht
    return th
