"""makani_b200 - B200-native (sm_100a) implementation of the SFNO spectral hot path of choutilin/makani:

    RealSHT -> SpectralConv (dhconv / diagonal) -> InverseRealSHT      (single GPU and h/w spatially sharded)

Python modules mirror the reference's operator API (torch-harmonics transforms + makani's SpectralConv);
the arithmetic runs in hand-written CUDA kernels behind a C ABI (include/makani_b200.h,
makani_b200/libmakani_b200.so).  There is no CPU path: using a module on a CPU tensor, or without the
built library, raises.
"""
from . import quadrature, legendre                      # host-side fp64 precompute (numpy only)
from .sht import RealSHT, InverseRealSHT
from .fft2 import RealFFT2, InverseRealFFT2
from . import losses
from .losses import GeometricH1Loss
from .spectral_convolution import SpectralConv, SpectralConvS2
from .instance_norm import FusedInstanceNorm2d, DistributedFusedInstanceNorm2d
from .block import fuse_block, fuse_model
from .pointwise_mlp import pointwise_linear, pointwise_mlp, fuse_pointwise
from . import distributed
from .compat import install_torch_harmonics_shim

__all__ = ["RealSHT", "InverseRealSHT", "RealFFT2", "InverseRealFFT2", "GeometricH1Loss", "losses", "SpectralConv",
           "SpectralConvS2", "FusedInstanceNorm2d", "DistributedFusedInstanceNorm2d", "fuse_block", "fuse_model",
           "fuse_pointwise", "pointwise_linear", "pointwise_mlp", "distributed", "quadrature", "legendre",
           "install_torch_harmonics_shim"]
__version__ = "0.1.0"
