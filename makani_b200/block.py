"""Rewire a ``FourierNeuralOperatorBlock`` (makani/models/networks/sfnonet.py:119-267) to the fused kernels -
SURVEY 8(f) rank 1.

The reference block is a plain ``nn.Module`` whose ``forward`` (sfnonet.py:239-267) composes

    filter -> norm0 [-> + inner_skip(residual)] -> act_layer0 -> mlp -> norm1 -> drop_path [-> + outer_skip(residual)]
           [-> act_layer1]

``fuse_block(block)`` replaces, in place and without touching parameter names (state dicts stay interchangeable):

* ``norm0`` + ``act_layer0``  by one ``FusedInstanceNorm2d(act="gelu")`` when there is no inner skip (the named
  configuration: ``inner_skip="none"``, sfnonet.py:395), or ``norm0`` + the inner-skip add by
  ``FusedInstanceNorm2d()(x, skip)`` when there is one;
* ``norm1`` + the outer-skip add by ``FusedInstanceNorm2d()(x, skip)`` when ``drop_path`` is the identity
  (``outer_skip="linear"``, sfnonet.py:396);
* the reference's ``SpectralConv`` inside ``block.filter`` is left alone - swap that class through the import shim
  (INTEGRATION.md section 1).

Only ``nn.InstanceNorm2d`` without running statistics (what sfnonet.py:379 builds) and exact ``nn.GELU`` are fused;
anything else is left as it is.  ``fuse_model(model)`` does this for every block of ``model.blocks``.
"""
from __future__ import annotations

import types

import torch.nn as nn

from .instance_norm import DistributedFusedInstanceNorm2d, FusedInstanceNorm2d
from .pointwise_mlp import fuse_pointwise

__all__ = ["fuse_block", "fuse_model", "fused_block_forward"]


def _fusable_norm(m) -> bool:
    if isinstance(m, FusedInstanceNorm2d):
        return False
    if isinstance(m, nn.InstanceNorm2d):
        return not m.track_running_stats
    # the reference's DistributedInstanceNorm2d (mpu/layer_norm.py:27-45): duck-typed, it is not importable here
    return type(m).__name__ == "DistributedInstanceNorm2d" and hasattr(m, "eps") and hasattr(m, "affine")


def _fused_from(norm, act: str) -> FusedInstanceNorm2d:
    distributed = type(norm).__name__ == "DistributedInstanceNorm2d"
    affine = bool(norm.affine)
    nfeat = norm.num_features if hasattr(norm, "num_features") else (norm.weight.numel() if affine else 0)
    if distributed:
        fused = DistributedFusedInstanceNorm2d(nfeat, eps=norm.eps, affine=affine, act=act)
    else:
        fused = FusedInstanceNorm2d(nfeat, eps=norm.eps, affine=affine, act=act)
    if affine:
        fused.weight = norm.weight            # the SAME Parameter objects: optimizer state and checkpoints carry over
        fused.bias = norm.bias
    return fused


def _exact_gelu(m) -> bool:
    return isinstance(m, nn.GELU) and getattr(m, "approximate", "none") == "none"


def fused_block_forward(self, x):
    """``FourierNeuralOperatorBlock.forward`` (sfnonet.py:239-267) with the fused norm calls."""
    x, residual = self.filter(x)

    if getattr(self, "_fused_inner_skip", False):
        x = self.norm0(x, self.inner_skip(residual))
    else:
        x = self.norm0(x)
        if hasattr(self, "inner_skip"):
            x = x + self.inner_skip(residual)

    if hasattr(self, "act_layer0"):
        x = self.act_layer0(x)

    if hasattr(self, "mlp"):
        x = self.mlp(x)

    if getattr(self, "_fused_outer_skip", False):
        x = self.norm1(x, self.outer_skip(residual))
    else:
        x = self.norm1(x)
        x = self.drop_path(x)
        if hasattr(self, "outer_skip"):
            x = x + self.outer_skip(residual)

    if hasattr(self, "act_layer1"):
        x = self.act_layer1(x)

    return x


def fuse_block(block: nn.Module, pointwise: bool = True) -> nn.Module:
    """In-place rewiring of one block; returns it.  Idempotent.  pointwise: also route the block's 1x1 convolutions
    (``mlp``, ``outer_skip`` / ``inner_skip``) through the GEMM-on-NCHW path of pointwise_mlp.py."""
    if pointwise:
        for name in ("mlp", "outer_skip", "inner_skip"):
            if hasattr(block, name):
                fuse_pointwise(getattr(block, name))
    if getattr(block, "_makani_b200_fused", False):
        return block
    for name in ("filter", "norm0", "norm1", "drop_path"):
        if not hasattr(block, name):
            raise TypeError(f"fuse_block: {type(block).__name__} has no attribute {name!r}; expected a "
                            "FourierNeuralOperatorBlock (sfnonet.py:119-267)")
    block._fused_inner_skip = False
    block._fused_outer_skip = False
    if _fusable_norm(block.norm0):
        if hasattr(block, "inner_skip"):
            block.norm0 = _fused_from(block.norm0, "none")
            block._fused_inner_skip = True
        elif hasattr(block, "act_layer0") and _exact_gelu(block.act_layer0):
            block.norm0 = _fused_from(block.norm0, "gelu")
            block.act_layer0 = nn.Identity()
        else:
            block.norm0 = _fused_from(block.norm0, "none")
    if _fusable_norm(block.norm1):
        block.norm1 = _fused_from(block.norm1, "none")
        if hasattr(block, "outer_skip") and isinstance(block.drop_path, nn.Identity):
            block._fused_outer_skip = True
    block.forward = types.MethodType(fused_block_forward, block)
    block._makani_b200_fused = True
    return block


def fuse_model(model: nn.Module, pointwise: bool = True) -> nn.Module:
    """``fuse_block`` on every element of ``model.blocks`` (SphericalFourierNeuralOperatorNet, sfnonet.py:386-431);
    with ``pointwise`` also the encoder, the decoder and the big-skip convolution (sfnonet.py:356-363, 451-468)."""
    if not hasattr(model, "blocks"):
        raise TypeError(f"fuse_model: {type(model).__name__} has no .blocks")
    for blk in model.blocks:
        fuse_block(blk, pointwise=pointwise)
    if pointwise:
        for name in ("encoder", "decoder", "residual_transform"):
            if hasattr(model, name):
                fuse_pointwise(getattr(model, name))
    return model
