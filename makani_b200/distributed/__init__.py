"""Spatially (h = latitude / w = longitude) sharded spherical harmonic transforms: drop-ins for
``torch_harmonics.distributed`` as the reference uses it (makani/models/networks/sfnonet.py:528-533,
makani/models/common/spectral_convolution.py:78-82, makani/mpu/layers.py:34-35,73-100).

    init(polar_group, azimuth_group)
    DistributedRealSHT / DistributedInverseRealSHT     (attrs lat_shapes, lon_shapes, l_shapes, m_shapes)
    distributed_transpose_polar / distributed_transpose_azimuth   (autograd Functions, .apply(x, dims, sizes))
    compute_split_shapes, split_tensor_along_dim

Choreography (the in-tree structural twin is makani/mpu/layers.py:67-102 / 134-169):

    [B, C, K/h, W/w] --a2a_w (C<->lon)--> [B, C/w, K/h, W]   longitude DFT, keep M
                     --a2a_w (m<->C)--->  [B, C, K/h, M/w]
                     --a2a_h (C<->lat)--> [B, C/h, K, M/w]   Legendre (table holds this rank's m-slab)
                     --a2a_h (l<->C)--->  [B, C, L/h, M/w]

The exchanges are NCCL all-to-alls over NVLink issued through torch.distributed (one process per GPU); the
local stages are the package's CUDA kernels.  The transposes themselves are device agnostic, so the
re-sharding logic is testable on CPU with gloo.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from .. import ops
from ..sht import InverseRealSHT, RealSHT

__all__ = ["init", "is_initialized", "polar_group", "azimuth_group", "polar_group_size", "azimuth_group_size",
           "polar_group_rank", "azimuth_group_rank", "compute_split_shapes", "split_tensor_along_dim",
           "distributed_transpose_polar", "distributed_transpose_azimuth", "DistributedRealSHT",
           "DistributedInverseRealSHT"]

_POLAR_GROUP = None
_AZIMUTH_GROUP = None
_IS_INITIALIZED = False


def init(polar_process_group, azimuth_process_group):
    """Either group may be None (that axis is not split)."""
    global _POLAR_GROUP, _AZIMUTH_GROUP, _IS_INITIALIZED
    _POLAR_GROUP = polar_process_group
    _AZIMUTH_GROUP = azimuth_process_group
    _IS_INITIALIZED = True


def is_initialized() -> bool:
    return _IS_INITIALIZED


def polar_group():
    return _POLAR_GROUP


def azimuth_group():
    return _AZIMUTH_GROUP


def _size(group) -> int:
    if group is None or not dist.is_initialized():
        return 1
    return dist.get_world_size(group=group)


def _rank(group) -> int:
    if group is None or not dist.is_initialized():
        return 0
    return dist.get_rank(group=group)


def polar_group_size() -> int:
    return _size(_POLAR_GROUP)


def azimuth_group_size() -> int:
    return _size(_AZIMUTH_GROUP)


def polar_group_rank() -> int:
    return _rank(_POLAR_GROUP)


def azimuth_group_rank() -> int:
    return _rank(_AZIMUTH_GROUP)


# ------------------------------------------------------------------------------------------------
def compute_split_shapes(size: int, num_chunks: int) -> List[int]:
    """Ceil-division chunks with a smaller last chunk (721/4 -> 181,181,181,178); floor chunks if the last
    one would be empty."""
    if num_chunks == 1:
        return [size]
    chunk = -(-size // num_chunks)
    last = max(0, size - chunk * (num_chunks - 1))
    if last == 0:
        chunk = size // num_chunks
        last = size - chunk * (num_chunks - 1)
    return [chunk] * (num_chunks - 1) + [last]


def split_tensor_along_dim(tensor: torch.Tensor, dim: int, num_chunks: int):
    if dim >= tensor.dim() or dim < -tensor.dim():
        raise ValueError(f"cannot split dim {dim} of a {tensor.dim()}-d tensor")
    if tensor.shape[dim] < num_chunks:
        raise ValueError(f"cannot split {tensor.shape[dim]} elements into {num_chunks} chunks")
    return torch.split(tensor, compute_split_shapes(tensor.shape[dim], num_chunks), dim=dim)


def _all_to_all(recv, send, group):
    """all-to-all of uneven chunks.  NCCL: one grouped send/recv over NVLink.  Backends without a list
    all-to-all (gloo, used by the CPU tests): pairwise isend/irecv."""
    if dist.get_backend(group) != "gloo":
        dist.all_to_all(recv, send, group=group)
        return
    rank = dist.get_rank(group=group)
    size = dist.get_world_size(group=group)
    cplx = [t.is_complex() for t in recv]
    rbuf = [torch.view_as_real(t) if c else t for t, c in zip(recv, cplx)]
    sbuf = [torch.view_as_real(t) if t.is_complex() else t for t in send]
    rbuf[rank].copy_(sbuf[rank])
    reqs = []
    for peer in range(size):
        if peer == rank:
            continue
        gpeer = dist.get_global_rank(group, peer) if group is not None else peer
        reqs.append(dist.isend(sbuf[peer].contiguous(), gpeer, group=group))
        reqs.append(dist.irecv(rbuf[peer], gpeer, group=group))
    for r in reqs:
        r.wait()


def _transpose(tensor: torch.Tensor, dim0: int, dim1: int, dim1_split_sizes, group):
    """Scatter dim0 over the group, gather dim1: returns (received chunks, the dim0 sizes that were sent)."""
    comm_size = dist.get_world_size(group=group)
    comm_rank = dist.get_rank(group=group)
    send = [t.contiguous() for t in split_tensor_along_dim(tensor, dim=dim0, num_chunks=comm_size)]
    send_shapes = [t.shape for t in send]
    shape = list(send_shapes[comm_rank])
    recv = []
    for n in dim1_split_sizes:
        shape[dim1] = n
        recv.append(torch.empty(shape, dtype=tensor.dtype, device=tensor.device))
    _all_to_all(recv, send, group)
    return recv, [s[dim0] for s in send_shapes]


def _transpose_fwd(group, ctx, x, dims, dim1_split_sizes):
    x = x.contiguous()
    chunks, dim0_split_sizes = _transpose(x, dims[0], dims[1], dim1_split_sizes, group)
    ctx.dims = dims
    ctx.dim0_split_sizes = dim0_split_sizes
    return torch.cat(chunks, dim=dims[1]).contiguous()


def _transpose_bwd(group, ctx, go):
    dims = ctx.dims
    chunks, _ = _transpose(go.contiguous(), dims[1], dims[0], ctx.dim0_split_sizes, group)
    return torch.cat(chunks, dim=dims[0]).contiguous(), None, None


class distributed_transpose_azimuth(torch.autograd.Function):
    """``.apply(x, (dim0, dim1), dim1_split_sizes)`` on the azimuth (w) group: dim0 is scattered, dim1 gathered."""

    @staticmethod
    def forward(ctx, x, dims, dim1_split_sizes):
        return _transpose_fwd(_AZIMUTH_GROUP, ctx, x, dims, dim1_split_sizes)

    @staticmethod
    def backward(ctx, go):
        return _transpose_bwd(_AZIMUTH_GROUP, ctx, go)


class distributed_transpose_polar(torch.autograd.Function):
    """``.apply(x, (dim0, dim1), dim1_split_sizes)`` on the polar (h) group: dim0 is scattered, dim1 gathered."""

    @staticmethod
    def forward(ctx, x, dims, dim1_split_sizes):
        return _transpose_fwd(_POLAR_GROUP, ctx, x, dims, dim1_split_sizes)

    @staticmethod
    def backward(ctx, go):
        return _transpose_bwd(_POLAR_GROUP, ctx, go)


# ------------------------------------------------------------------------------------------------
# local compute stages on complex64 tensors (between the all-to-alls)
class _DFTStage(torch.autograd.Function):
    """x real [B, C, Kloc, W] -> 2*pi*rfft(x, norm="forward")[..., :M] complex64 [B, C, Kloc, M]."""

    @staticmethod
    def forward(ctx, x, mod):
        B, Cn, K, W = x.shape
        xm = ops.dft_fwd(x, mod.mmax, mod._scale(x.device, "fwd"))
        ctx.mod, ctx.shape, ctx.in_dtype = mod, (B, Cn, K, W), x.dtype
        return ops.xm_to_c64(xm, B, Cn, K, mod.mmax)

    @staticmethod
    def backward(ctx, g):
        mod = ctx.mod
        B, Cn, K, W = ctx.shape
        gxm = ops.c64_to_xm(g.to(torch.complex64))
        return ops.dft_inv(gxm, B, Cn, K, W, mod.mmax, mod._scale(g.device, "rfft_adj"), ctx.in_dtype), None


class _IDFTStage(torch.autograd.Function):
    """complex64 [B, C, Kloc, M] -> irfft(n=W, norm="forward") fp32 [B, C, Kloc, W]."""

    @staticmethod
    def forward(ctx, c, mod):
        B, Cn, K, M = c.shape
        xm = ops.c64_to_xm(c)
        ctx.mod, ctx.shape = mod, (B, Cn, K, M)
        return ops.dft_inv(xm, B, Cn, K, mod.nlon, M, mod._scale(c.device, "one"), torch.float32)

    @staticmethod
    def backward(ctx, g):
        mod = ctx.mod
        B, Cn, K, M = ctx.shape
        gxm = ops.dft_fwd(g.contiguous(), M, mod._scale(g.device, "irfft_adj"))
        return ops.xm_to_c64(gxm, B, Cn, K, M), None


class _LegendreStage(torch.autograd.Function):
    """analysis (synthesis=False): [B, C, K, Mloc] -> [B, C, L, Mloc];  synthesis: the reverse.  complex64."""

    @staticmethod
    def forward(ctx, c, mod, synthesis):
        plan = mod._plan(c.device)
        B, Cn = c.shape[0], c.shape[1]
        ctx.mod, ctx.synthesis = mod, synthesis
        if synthesis:
            xm = ops.legendre_synthesis(plan, ops.c64_to_sp(c), B, Cn)
            return ops.xm_to_c64(xm, B, Cn, mod.nlat, plan.mmax)
        sp = ops.legendre_analysis(plan, ops.c64_to_xm(c), B, Cn)
        return ops.sp_to_c64(sp, B, Cn, mod.lmax, plan.mmax, 0, plan.m_offset)

    @staticmethod
    def backward(ctx, g):
        mod = ctx.mod
        plan = mod._plan(g.device)
        B, Cn = g.shape[0], g.shape[1]
        g = g.to(torch.complex64)
        if ctx.synthesis:
            gsp = ops.legendre_analysis(plan, ops.c64_to_xm(g), B, Cn)
            return ops.sp_to_c64(gsp, B, Cn, mod.lmax, plan.mmax, 0, plan.m_offset), None, None
        gxm = ops.legendre_synthesis(plan, ops.c64_to_sp(g), B, Cn)
        return ops.xm_to_c64(gxm, B, Cn, mod.nlat, plan.mmax), None, None


class _DistributedMixin:
    def _setup_shards(self):
        self.comm_size_polar = polar_group_size()
        self.comm_rank_polar = polar_group_rank()
        self.comm_size_azimuth = azimuth_group_size()
        self.comm_rank_azimuth = azimuth_group_rank()
        self.lat_shapes = compute_split_shapes(self.nlat, self.comm_size_polar)
        self.lon_shapes = compute_split_shapes(self.nlon, self.comm_size_azimuth)
        self.l_shapes = compute_split_shapes(self.lmax, self.comm_size_polar)
        self.m_shapes = compute_split_shapes(self.mmax, self.comm_size_azimuth)

    def _table_slice(self):
        m0 = sum(self.m_shapes[: self.comm_rank_azimuth])
        m1 = m0 + self.m_shapes[self.comm_rank_azimuth]
        return m0, np.ascontiguousarray(self._table64()[m0:m1])

    # local stages; tests replace these with the CPU oracle to exercise the choreography without a GPU
    def _dft_stage(self, x):
        return _DFTStage.apply(x, self)

    def _idft_stage(self, c):
        return _IDFTStage.apply(c, self)

    def _legendre_stage(self, c, synthesis):
        return _LegendreStage.apply(c, self, synthesis)


class DistributedRealSHT(_DistributedMixin, RealSHT):
    """Forward SHT of a field sharded as [B, C, nlat/h, nlon/w] -> coefficients [B, C, lmax/h, mmax/w]."""

    def __init__(self, nlat, nlon, lmax=None, mmax=None, grid="lobatto", norm="ortho", csphase=True):
        RealSHT.__init__(self, nlat, nlon, lmax, mmax, grid, norm, csphase)
        self._setup_shards()
        m0, tab = self._table_slice()
        self.register_buffer("weights", torch.from_numpy(tab), persistent=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() != 4:
            raise ValueError(f"Expected a 4-d tensor [B, C, H_local, W_local], got {x.dim()} dims")
        num_chans = x.shape[1]
        if self.comm_size_azimuth > 1:
            x = distributed_transpose_azimuth.apply(x, (1, -1), self.lon_shapes)
        x = self._dft_stage(x.contiguous())
        if self.comm_size_azimuth > 1:
            chan_shapes = compute_split_shapes(num_chans, self.comm_size_azimuth)
            x = distributed_transpose_azimuth.apply(x, (-1, 1), chan_shapes)
        if self.comm_size_polar > 1:
            x = distributed_transpose_polar.apply(x, (1, -2), self.lat_shapes)
        x = self._legendre_stage(x.contiguous(), False)
        if self.comm_size_polar > 1:
            chan_shapes = compute_split_shapes(num_chans, self.comm_size_polar)
            x = distributed_transpose_polar.apply(x, (-2, 1), chan_shapes)
        return x


class DistributedInverseRealSHT(_DistributedMixin, InverseRealSHT):
    """Inverse SHT of coefficients sharded as [B, C, lmax/h, mmax/w] -> field [B, C, nlat/h, nlon/w]."""

    def __init__(self, nlat, nlon, lmax=None, mmax=None, grid="lobatto", norm="ortho", csphase=True):
        InverseRealSHT.__init__(self, nlat, nlon, lmax, mmax, grid, norm, csphase)
        self._setup_shards()
        m0, tab = self._table_slice()
        self.register_buffer("pct", torch.from_numpy(tab), persistent=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() != 4:
            raise ValueError(f"Expected a 4-d tensor [B, C, L_local, M_local], got {x.dim()} dims")
        num_chans = x.shape[1]
        if self.comm_size_polar > 1:
            x = distributed_transpose_polar.apply(x, (1, -2), self.l_shapes)
        x = self._legendre_stage(x.contiguous(), True)
        if self.comm_size_polar > 1:
            chan_shapes = compute_split_shapes(num_chans, self.comm_size_polar)
            x = distributed_transpose_polar.apply(x, (-2, 1), chan_shapes)
        if self.comm_size_azimuth > 1:
            x = distributed_transpose_azimuth.apply(x, (1, -1), self.m_shapes)
        x = self._idft_stage(x.contiguous())
        if self.comm_size_azimuth > 1:
            chan_shapes = compute_split_shapes(num_chans, self.comm_size_azimuth)
            x = distributed_transpose_azimuth.apply(x, (-1, 1), chan_shapes)
        return x
