"""Fused instance norm (+ GELU, + skip add) for the layers around the spectral filter - SURVEY 8(f) rank 1.

In ``FourierNeuralOperatorBlock.forward`` (makani/models/networks/sfnonet.py:239-267) the named configuration runs

    x, residual = self.filter(x)
    x = self.act_layer0(self.norm0(x))                  # nn.InstanceNorm2d(C, eps=1e-6, affine=True), nn.GELU()
    x = self.mlp(x)
    x = self.norm1(x) + self.outer_skip(residual)

``FusedInstanceNorm2d`` computes ``act(norm(x)) (+ skip)`` with two bandwidth kernels forward (statistics, apply) and
two backward (csrc/pointwise.cu) instead of the five / eight full-tensor passes of the separate torch modules.  Same
parameters as ``nn.InstanceNorm2d`` (``weight``, ``bias``; no running statistics), so state dicts are interchangeable;
INTEGRATION.md shows the patch of the block.

``DistributedFusedInstanceNorm2d`` is the spatially sharded variant with the semantics of the reference's
``DistributedInstanceNorm2d`` (makani/mpu/layer_norm.py:27-114): every rank computes the moments of its slab, the
(mean, variance, count) triples are all-gathered over the "spatial" group and merged (pairwise update of Chan et al.,
the formula the reference applies sequentially), backward all-reduces the two per-plane sums.

CUDA only: no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from ._nvtx import annotate
from . import _ext
from .ops import _need_cuda, _ptr, _stream, dtype_code

__all__ = ["FusedInstanceNorm2d", "DistributedFusedInstanceNorm2d", "instance_norm_act", "merge_moments"]

_ACTS = {"none": _ext.ACT_NONE, None: _ext.ACT_NONE, "gelu": _ext.ACT_GELU}


def _opt(t: Optional[torch.Tensor]):
    return C.c_void_p(0) if t is None else _ptr(t)


def merge_moments(mean: torch.Tensor, var: torch.Tensor, count: torch.Tensor):
    """Merge per-rank moments along the last dimension: mean / var [..., R] (biased variances), count [R] or [..., R].
    Pairwise update of Chan, Golub & LeVeque folded left to right, as layer_norm.py:66-80 does.  Pure tensor
    arithmetic (device agnostic).  Returns (mean, biased variance, count)."""
    count = count.to(mean.dtype).expand_as(mean)
    m, m2, n = mean[..., 0], var[..., 0] * count[..., 0], count[..., 0]
    for i in range(1, mean.shape[-1]):
        ni = count[..., i]
        delta = mean[..., i] - m
        tot = n + ni
        m2 = m2 + var[..., i] * ni + delta * delta * n * ni / tot
        m = m + delta * ni / tot
        n = tot
    return m, m2 / n, n


def _stats(x: torch.Tensor, eps: float, want_moments: bool):
    B, Cn, H, W = x.shape
    planes, hw = B * Cn, H * W
    ws = torch.empty(planes * _ext.INSTNORM_SEGMENTS * 2, dtype=torch.float32, device=x.device)
    stats = torch.empty(planes, 2, dtype=torch.float32, device=x.device)
    mv = torch.empty(planes, 2, dtype=torch.float32, device=x.device) if want_moments else None
    _ext.check(_ext.load().sfno_instnorm_stats(_ptr(x), dtype_code(x.dtype), planes, hw, float(eps), _ptr(ws),
                                               _opt(mv), _ptr(stats), _stream()), "sfno_instnorm_stats")
    return stats, mv


class _InstanceNormAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, skip, eps, act, group):
        _need_cuda(x, "x")
        if x.dim() != 4:
            raise ValueError(f"expected [B, C, H, W], got {tuple(x.shape)}")
        x = x.contiguous()
        B, Cn, H, W = x.shape
        planes, hw = B * Cn, H * W
        dt = dtype_code(x.dtype)
        if skip is not None:
            if act != _ext.ACT_NONE:
                raise ValueError("a skip input is only supported with act='none' (norm1(x) + outer_skip(residual))")
            if skip.shape != x.shape or skip.dtype != x.dtype:
                raise ValueError(f"skip must match x: {tuple(skip.shape)} {skip.dtype} vs {tuple(x.shape)} {x.dtype}")
            skip = skip.contiguous()
        w32 = None if weight is None else weight.detach().to(torch.float32).contiguous()
        b32 = None if bias is None else bias.detach().to(torch.float32).contiguous()

        sharded = group is not None and dist.get_world_size(group=group) > 1
        stats, mv = _stats(x, eps, want_moments=sharded)
        ntot = None
        if sharded:
            # merge the ranks' (mean, var, count): tiny [planes] tensors, one all-gather
            R = dist.get_world_size(group=group)
            mine = torch.cat([mv, torch.full((planes, 1), float(hw), dtype=torch.float32, device=x.device)], dim=1)
            allm = [torch.empty_like(mine) for _ in range(R)]
            dist.all_gather(allm, mine, group=group)
            allm = torch.stack(allm, dim=-1)                                   # [planes, 3, R]
            mean, var, n = merge_moments(allm[:, 0], allm[:, 1], allm[:, 2])
            stats = torch.stack([mean, torch.rsqrt(var + eps)], dim=1).contiguous()
            # elements of the whole instance, per plane, kept ON THE DEVICE (no host read: the reference captures the
            # step in a CUDA graph); the backward pass folds it into the per-plane sums
            ntot = n.contiguous()
        out = torch.empty_like(x)
        _ext.check(_ext.load().sfno_instnorm_apply(_ptr(x), _opt(skip), dt, planes, Cn, hw, _ptr(stats), _opt(w32),
                                                   _opt(b32), act, _ptr(out), _stream()), "sfno_instnorm_apply")
        tensors = [x, stats] + [t for t in (w32, b32, ntot) if t is not None]
        ctx.save_for_backward(*tensors)
        ctx.cfg = (act, group if sharded else None, skip is not None, w32 is not None, b32 is not None,
                   None if weight is None else weight.dtype)
        return out

    @staticmethod
    def backward(ctx, g):
        act, group, has_skip, has_w, has_b, wdtype = ctx.cfg
        saved = list(ctx.saved_tensors)
        x, stats = saved[0], saved[1]
        w32 = saved[2] if has_w else None
        b32 = saved[2 + int(has_w)] if has_b else None
        ntot = saved[2 + int(has_w) + int(has_b)] if group is not None else None
        B, Cn, H, W = x.shape
        planes, hw = B * Cn, H * W
        dt = dtype_code(x.dtype)
        g = g.contiguous()
        lib = _ext.load()
        part = torch.empty(planes, _ext.INSTNORM_SEGMENTS, 2, dtype=torch.float32, device=x.device)
        _ext.check(lib.sfno_instnorm_bwd_stats(_ptr(g), _ptr(x), dt, planes, Cn, hw, _ptr(stats), _opt(w32), _opt(b32),
                                               act, _ptr(part), _stream()), "sfno_instnorm_bwd_stats")
        local = part.sum(dim=1)                                                 # [planes, 2]: (sum dz, sum dz * xhat)
        bsum = local
        if group is not None:
            bsum = local.clone()
            dist.all_reduce(bsum, group=group)
            # the kernel divides the sums by its `hw_total` argument; passing the local size and pre-scaling the sums
            # by hw / n_total (device arithmetic) gives sums / n_total without reading n_total back to the host
            bsum = (bsum * (float(hw) / ntot).unsqueeze(1)).contiguous()
        gx = gw = gb = None
        if ctx.needs_input_grad[0]:
            gx = torch.empty_like(x)
            _ext.check(lib.sfno_instnorm_bwd_apply(_ptr(g), _ptr(x), dt, planes, Cn, hw, hw, _ptr(stats),
                                                   _opt(w32), _opt(b32), _ptr(bsum), act, _ptr(gx), _stream()),
                       "sfno_instnorm_bwd_apply")
        # affine-parameter gradients: sums over the batch of this rank's per-plane sums.  For a sharded instance they
        # are partial (this rank's slab), exactly like the reference's: the parameters are marked
        # is_shared_mp = ["spatial"] (layer_norm.py:44-45) and the gradient hooks sum them over that group.
        per = local.view(B, Cn, 2)
        if has_w and ctx.needs_input_grad[1]:
            gw = per[..., 1].sum(dim=0).to(wdtype)
        if has_b and ctx.needs_input_grad[2]:
            gb = per[..., 0].sum(dim=0).to(wdtype)
        gskip = g if (has_skip and ctx.needs_input_grad[3]) else None
        return gx, gw, gb, gskip, None, None, None


def instance_norm_act(x, weight=None, bias=None, skip=None, eps: float = 1e-5, act: str = "none", group=None):
    """``act(instance_norm(x) * weight + bias) (+ skip)`` on a CUDA tensor [B, C, H, W] (fp32 / bf16 / fp16)."""
    if act not in _ACTS:
        raise ValueError(f"unknown activation {act!r}; supported: 'none', 'gelu'")
    if skip is not None and skip.dtype != x.dtype:
        # mixed dtypes (e.g. an fp32 field plus a bf16 autocast conv output): promote like ``x + skip`` would
        dt = torch.promote_types(x.dtype, skip.dtype)
        x, skip = x.to(dt), skip.to(dt)
    return _InstanceNormAct.apply(x, weight, bias, skip, float(eps), _ACTS[act], group)


class FusedInstanceNorm2d(nn.Module):
    """``nn.InstanceNorm2d(num_features, eps, affine)`` (no running statistics) fused with an optional activation
    (``act="gelu"``: ``act_layer0(norm0(x))``) or skip add (``forward(x, skip)``: ``norm1(x) + outer_skip(residual)``)."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=False, track_running_stats=False, act="none",
                 device=None, dtype=None):
        super().__init__()
        if track_running_stats:
            raise NotImplementedError("running statistics are not supported (the reference passes False, sfnonet.py:379)")
        if act not in _ACTS:
            raise ValueError(f"unknown activation {act!r}; supported: 'none', 'gelu'")
        self.num_features, self.eps, self.affine, self.act = num_features, eps, affine, act or "none"
        if affine:
            self.weight = nn.Parameter(torch.ones(num_features, device=device, dtype=dtype))
            self.bias = nn.Parameter(torch.zeros(num_features, device=device, dtype=dtype))
        else:
            self.register_parameter("weight", None)
            self.register_parameter("bias", None)

    def _group(self):
        return None

    @annotate("FusedInstanceNorm2d.forward")
    def forward(self, x, skip=None):
        if self.num_features and x.shape[1] != self.num_features:      # 0 / None: unknown (non-affine norm adopted by fuse_block)
            raise ValueError(f"expected {self.num_features} channels, got {x.shape[1]}")
        return instance_norm_act(x, self.weight, self.bias, skip, self.eps, self.act, self._group())

    def extra_repr(self):
        return f"{self.num_features}, eps={self.eps}, affine={self.affine}, act={self.act}"


class DistributedFusedInstanceNorm2d(FusedInstanceNorm2d):
    """Spatially sharded instance norm: same constructor as ``DistributedInstanceNorm2d`` (layer_norm.py:31-45) plus
    ``act``; statistics over the "spatial" group of ``makani_b200.comm`` (or an explicit ``group``)."""

    def __init__(self, num_features, eps=1e-05, affine=False, device=None, dtype=None, act="none", group=None):
        super().__init__(num_features, eps=eps, affine=affine, act=act, device=device, dtype=dtype)
        self._explicit_group = group
        if self.affine:
            self.weight.is_shared_mp = ["spatial"]
            self.bias.is_shared_mp = ["spatial"]

    def _group(self):
        if self._explicit_group is not None:
            return self._explicit_group
        from . import comm
        return comm.get_group("spatial") if comm.get_size("spatial") > 1 else None
