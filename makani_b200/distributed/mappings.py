"""The four parallel-region mappings the reference imports from ``modulus.distributed.mappings`` (copy / reduce /
scatter / gather with their adjoint collectives in backward), over ``torch.distributed`` and the group registry of
``makani_b200.comm``.  SURVEY 8(f) rank 2.

Call sites in the reference: ``mpu/layer_norm.py:50,100-101`` (instance-norm statistics over "spatial"),
``mpu/layers.py:265-267,442-445`` (feature-parallel MLP), ``sfnonet.py:605,635``, ``utils/losses.py:153-155``.
Signatures follow those call sites: the group is named, not passed::

    copy_to_parallel_region(x, "spatial")                 identity forward, all-reduce backward
    reduce_from_parallel_region(x, "spatial")             all-reduce forward, identity backward
    scatter_to_parallel_region(x, dim, "fin")             keep this rank's chunk forward, all-gather backward
    gather_from_parallel_region(x, dim, shapes, "h")      all-gather (uneven ``shapes`` along dim) forward, chunk backward

All of them are no-ops when the named group has one rank.  Device agnostic (NCCL on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from .. import comm
from . import compute_split_shapes

__all__ = ["copy_to_parallel_region", "reduce_from_parallel_region", "scatter_to_parallel_region",
           "gather_from_parallel_region"]


def _all_reduce(x: torch.Tensor, group) -> torch.Tensor:
    x = x.contiguous().clone()
    buf = torch.view_as_real(x) if x.is_complex() else x
    dist.all_reduce(buf, group=group)
    return x


def _split(x: torch.Tensor, dim: int, group, shapes: Optional[List[int]] = None) -> torch.Tensor:
    size, rank = dist.get_world_size(group=group), dist.get_rank(group=group)
    sizes = list(shapes) if shapes is not None else compute_split_shapes(x.shape[dim], size)
    return torch.split(x, sizes, dim=dim)[rank].contiguous()


def _all_gather_v(x: torch.Tensor, dim: int, shapes: Optional[List[int]], group) -> torch.Tensor:
    size, rank = dist.get_world_size(group=group), dist.get_rank(group=group)
    dim = dim % x.dim()
    x = x.contiguous()
    sizes = list(shapes) if shapes is not None else [x.shape[dim]] * size
    assert sizes[rank] == x.shape[dim], (sizes, rank, x.shape)
    if len(set(sizes)) == 1:
        out = [torch.empty_like(x) for _ in range(size)]
        bufs = [torch.view_as_real(t) if t.is_complex() else t for t in out]
        dist.all_gather(bufs, torch.view_as_real(x) if x.is_complex() else x, group=group)
        return torch.cat(out, dim=dim)
    # uneven chunks: pad to the largest along `dim`, gather, trim
    big = max(sizes)
    shape = list(x.shape)
    shape[dim] = big
    xp = x.new_zeros(shape)
    xp.narrow(dim, 0, x.shape[dim]).copy_(x)
    out = [torch.empty_like(xp) for _ in range(size)]
    bufs = [torch.view_as_real(t) if t.is_complex() else t for t in out]
    dist.all_gather(bufs, torch.view_as_real(xp) if xp.is_complex() else xp, group=group)
    return torch.cat([t.narrow(dim, 0, s) for t, s in zip(out, sizes)], dim=dim)


class _Copy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return _all_reduce(g, ctx.group), None


class _Reduce(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        return _all_reduce(x, group)

    @staticmethod
    def backward(ctx, g):
        return g, None


class _Scatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim, group):
        ctx.dim, ctx.group = dim, group
        ctx.sizes = compute_split_shapes(x.shape[dim], dist.get_world_size(group=group))
        return _split(x, dim, group)

    @staticmethod
    def backward(ctx, g):
        return _all_gather_v(g, ctx.dim, ctx.sizes, ctx.group), None, None


class _Gather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim, shapes, group):
        size = dist.get_world_size(group=group)
        ctx.dim, ctx.group = dim, group
        ctx.sizes = list(shapes) if shapes is not None else [x.shape[dim]] * size
        return _all_gather_v(x, dim, ctx.sizes, group)

    @staticmethod
    def backward(ctx, g):
        return _split(g, ctx.dim, ctx.group, ctx.sizes), None, None, None


def copy_to_parallel_region(x: torch.Tensor, group_name: str) -> torch.Tensor:
    if comm.get_size(group_name) == 1:
        return x
    return _Copy.apply(x, comm.get_group(group_name))


def reduce_from_parallel_region(x: torch.Tensor, group_name: str) -> torch.Tensor:
    if comm.get_size(group_name) == 1:
        return x
    return _Reduce.apply(x, comm.get_group(group_name))


def scatter_to_parallel_region(x: torch.Tensor, dim: int, group_name: str) -> torch.Tensor:
    if comm.get_size(group_name) == 1:
        return x
    return _Scatter.apply(x, dim, comm.get_group(group_name))


def gather_from_parallel_region(x: torch.Tensor, dim: int, shapes, group_name: str) -> torch.Tensor:
    if comm.get_size(group_name) == 1:
        return x
    return _Gather.apply(x, dim, shapes, comm.get_group(group_name))
