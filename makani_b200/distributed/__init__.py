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

h-sharded transforms (w = 1, the configuration the reference benchmarks: ``h_parallel_size=4``) on one NVLink node
skip the complex64 round trips and the pack / all-to-all / unpack: each of their two exchanges is a single kernel that
stores rows of the internal layout straight into the owner's buffer in peer memory (``p2p.py``, ``_PolarExchange``).
``p2p.status()`` reports which path is active; ``MAKANI_B200_P2P=0`` forces the all-to-all path.
"""
from __future__ import annotations

import ctypes as ct
import os
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from .. import _ext, ops
from .._nvtx import annotate
from ..sht import InverseRealSHT, RealSHT
from . import p2p

__all__ = ["init", "is_initialized", "polar_group", "azimuth_group", "polar_group_size", "azimuth_group_size",
           "polar_group_rank", "azimuth_group_rank", "compute_split_shapes", "split_tensor_along_dim",
           "distributed_transpose_polar", "distributed_transpose_azimuth", "DistributedRealSHT",
           "DistributedInverseRealSHT"]

# Which transforms fuse the longitude DFT with the latitude <-> channel exchange (tensor stores / loads on peer memory,
# csrc/dft_p2.cuh PeerMaps) instead of running the DFT locally and moving rows with sfno_rows_scatter.  The fused
# kernels touch peer memory in the 64-byte pieces of their staging tiles (16 latitudes x 4 bytes per bin), the row
# scatter in whole latitude slabs (0.2 - 1.5 KB), and NVLink is far less efficient at 64 bytes.  Measured, C = 384,
# SHT / inverse SHT in ms (fused vs DFT + scatter):  721x1440 h = 2: 0.63 / 0.70 vs 0.70 / 0.78;  721x1440 h = 4:
# 0.66 / 0.65 vs 0.61 / 0.61;  240x480 h = 2: 0.30 / 0.32 vs 0.28 / 0.30.  So "auto" fuses only nlon = 1440 on two
# ranks (where the three FFTs per row hide the slower link traffic); "all" / "0" force either side.
_FUSED_DFT = os.environ.get("MAKANI_B200_P2P_FUSED_DFT", "auto")
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
            xm = ops.legendre_synthesis(plan, ops.c64_to_sp(c, 0, plan.m_offset), B, Cn)
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
        gxm = ops.legendre_synthesis(plan, ops.c64_to_sp(g, 0, plan.m_offset), B, Cn)
        return ops.xm_to_c64(gxm, B, Cn, mod.nlat, plan.mmax), None, None


# ------------------------------------------------------------------------------------------------
# h-sharded transforms over peer memory: the two transpositions of each transform are single kernels that move rows of
# the internal layouts (XM, SP) straight into the owner's buffers over NVLink (see p2p.py); w must be 1.
def _offsets(shapes):
    out, acc = [], 0
    for n in shapes:
        out.append(acc)
        acc += n
    return out


def _rup4(n):
    return (n + 3) // 4 * 4


def _rup8(n):
    return (n + 7) // 8 * 8


class _PolarExchange:
    """The four row scatters between  XM_loc [M][2B][Cp][Kp(lat slab)]  <->  XM_ch [M][2B][Cp(ch slab)][Kp(nlat)]
    and  SP_ch [L][M][2B][Cp(ch slab)]  <->  SP_l [L slab][M][2B][Cp]."""

    def __init__(self, mod, B, C, device):
        self.group = polar_group()
        self.me, self.h = mod.comm_rank_polar, mod.comm_size_polar
        self.B, self.C, self.M, self.L, self.J = B, C, mod.mmax, mod.lmax, 2 * B
        self.lat, self.lat0 = mod.lat_shapes, _offsets(mod.lat_shapes)
        self.ls, self.l0 = mod.l_shapes, _offsets(mod.l_shapes)
        self.cs = compute_split_shapes(C, self.h)
        self.c0 = _offsets(self.cs)
        # XM_ch keeps every rank's latitude slab at a 32-byte aligned offset (slabs padded to multiples of 8 latitudes;
        # the Legendre table of the peer path, _peer_plan, has zero columns at the pad positions), so the rows that
        # cross NVLink are 16-byte vector stores into whole sectors
        self.latp0 = _offsets([_rup8(k) for k in self.lat])
        self.Cp, self.Kpg = _rup4(C), sum(_rup8(k) for k in self.lat)
        self.device = device
        self._plans = {}

    @staticmethod
    def of(mod, B, C, device):
        cache = mod.__dict__.setdefault("_exchanges", {})
        k = (B, C, device.index)
        ex = cache.get(k)
        if ex is None:
            ex = _PolarExchange(mod, B, C, device)
            cache[k] = ex
        return ex

    def _run(self, key, nfloats, src, make_desc, out_floats):
        plan = self._plans.get(key)
        if plan is None:
            # receive buffers and the (static) row descriptors for either of them, built once per layout
            pb = p2p.buffers((key, self.B, self.C, self.M, self.L, self.Kpg), nfloats, self.group, self.device)
            arrs = []
            for b in range(2):
                descs = [d for d in (make_desc(p, pb.peer_ptrs[b][p]) for p in range(self.h)) if d is not None]
                arrs.append(p2p.desc_array(descs) if descs else None)
            plan = (pb, arrs, p2p.fence(self.group, self.device))
            self._plans[key] = plan
        pb, arrs, fence = plan
        b = pb.next()
        fence.pre_write()
        if arrs[b] is not None:
            p2p.scatter(src, arrs[b], boxed=key in ("xm_ch", "xm_lat"))
        fence()
        return pb.local[b][:out_floats]

    @staticmethod
    def _desc(dst, src_off, dst_off, n, length, ss, ds, tri=-1):
        if min(n) <= 0 or length <= 0:
            return None
        return _ext.CopyDesc(dst, src_off, dst_off, n[0], n[1], n[2], length, ss[0], ss[1], ss[2], ds[0], ds[1], ds[2], tri)

    def lat_to_chan(self, xm_loc):
        """XM of my latitude slab, all channels -> XM of all latitudes, my channel slab."""
        M, J, Cp, Kpg, me = self.M, self.J, self.Cp, self.Kpg, self.me
        Kl = _rup4(self.lat[me])

        def mk(p, dst):
            Cpp = _rup4(self.cs[p])
            return self._desc(dst, self.c0[p] * Kl, self.latp0[me], (M, J, self.cs[p]), Kl,
                              (J * Cp * Kl, Cp * Kl, Kl), (J * Cpp * Kpg, Cpp * Kpg, Kpg))
        n = M * J * max(_rup4(c) for c in self.cs) * Kpg
        return self._run("xm_ch", n, xm_loc, mk, M * J * _rup4(self.cs[me]) * Kpg)

    def chan_to_lat(self, xm_ch):
        """XM of all latitudes, my channel slab -> XM of my latitude slab, all channels."""
        M, J, Cp, Kpg, me = self.M, self.J, self.Cp, self.Kpg, self.me
        Cpm = _rup4(self.cs[me])

        def mk(p, dst):
            Kpp = _rup4(self.lat[p])
            return self._desc(dst, self.latp0[p], self.c0[me] * Kpp, (M, J, self.cs[me]), Kpp,
                              (J * Cpm * Kpg, Cpm * Kpg, Kpg), (J * Cp * Kpp, Cp * Kpp, Kpp))
        n = M * J * Cp * max(_rup4(k) for k in self.lat)
        return self._run("xm_lat", n, xm_ch, mk, M * J * Cp * _rup4(self.lat[me]))

    # ---- longitude DFT fused with the exchange (csrc/dft_p2.cuh, PeerMaps): nlon = 480 / 1440 only
    def dft_fused(self, mod) -> bool:
        if _FUSED_DFT == "0" or (_FUSED_DFT != "all" and not (mod.nlon == 1440 and self.h == 2)):
            return False
        return bool(_ext.load().sfno_dft_peer_supported(mod.nlon, mod.mmax))

    def _xm_ch_buffers(self):
        M, J = self.M, self.J
        n = M * J * max(_rup4(c) for c in self.cs) * self.Kpg
        plan = self._plans.get("xm_ch_fused")
        if plan is None:
            pb = p2p.buffers(("xm_ch", self.B, self.C, self.M, self.L, self.Kpg), n, self.group, self.device)
            ptrs = [(ct.c_void_p * self.h)(*pb.peer_ptrs[b]) for b in range(2)]
            shapes = (ct.c_int * self.h)(*self.cs)
            plan = (pb, ptrs, shapes, p2p.fence(self.group, self.device))
            self._plans["xm_ch_fused"] = plan
        return plan

    def dft_to_chan(self, x, mscale):
        """x [B, C, K_loc, W] -> XM of all latitudes, my channel slab: the DFT kernel's tensor stores write every tile
        into the buffer of the rank that owns its channel."""
        pb, ptrs, shapes, fence = self._xm_ch_buffers()
        b = pb.next()
        me = self.me
        fence.pre_write()
        _ext.check(_ext.load().sfno_dft_fwd_peer(
            ct.c_void_p(x.data_ptr()), ops.dtype_code(x.dtype), self.B, self.C, self.lat[me], x.shape[-1], self.M,
            ct.c_void_p(mscale.data_ptr()), ct.cast(ptrs[b], ct.c_void_p), ct.cast(shapes, ct.c_void_p), self.h, self.Kpg,
            self.latp0[me], _rup8(self.lat[me]), 1, ct.c_void_p(torch.cuda.current_stream().cuda_stream)),
            "sfno_dft_fwd_peer")
        p2p._STATUS["exchanges"] += 1
        fence()
        return pb.local[b][: self.M * self.J * _rup4(self.cs[me]) * self.Kpg]

    def chan_to_field(self, plan, sp_ch, nlon, mscale, dtype):
        """SP of all degrees, my channel slab -> field [B, C, K_loc, W]: Legendre synthesis into my XM_ch buffer, then
        every rank's inverse DFT kernel loads its tiles from the owners' buffers."""
        pb, ptrs, shapes, fence = self._xm_ch_buffers()
        b = pb.next()
        me = self.me
        fence.pre_write()
        ops.legendre_synthesis(plan, sp_ch, self.B, self.cs[me],
                               out=pb.local[b][: self.M * self.J * _rup4(self.cs[me]) * self.Kpg])
        fence()
        y = torch.empty((self.B, self.C, self.lat[me], nlon), dtype=dtype, device=self.device)
        _ext.check(_ext.load().sfno_dft_inv_peer(
            ct.cast(ptrs[b], ct.c_void_p), ct.cast(shapes, ct.c_void_p), self.h, self.Kpg, self.latp0[me],
            _rup8(self.lat[me]), self.B, self.C, self.lat[me], nlon, self.M, self.M, ct.c_void_p(mscale.data_ptr()),
            ct.c_void_p(y.data_ptr()), ops.dtype_code(dtype), ct.c_void_p(torch.cuda.current_stream().cuda_stream)),
            "sfno_dft_inv_peer")
        p2p._STATUS["exchanges"] += 1
        return y

    def chan_to_l(self, sp_ch):
        """SP of all degrees, my channel slab -> SP of my degree slab, all channels."""
        M, J, Cp, me = self.M, self.J, self.Cp, self.me
        Cpm = _rup4(self.cs[me])

        def mk(p, dst):
            return self._desc(dst, self.l0[p] * M * J * Cpm, self.c0[me], (self.ls[p], M, J), self.cs[me],
                              (M * J * Cpm, J * Cpm, Cpm), (M * J * Cp, J * Cp, Cp), tri=self.l0[p])
        n = max(self.ls) * M * J * Cp
        return self._run("sp_l", n, sp_ch, mk, self.ls[me] * M * J * Cp)

    def l_to_chan(self, sp_l):
        """SP of my degree slab, all channels -> SP of all degrees, my channel slab."""
        M, J, Cp, me = self.M, self.J, self.Cp, self.me

        def mk(p, dst):
            Cpp = _rup4(self.cs[p])
            return self._desc(dst, self.c0[p], self.l0[me] * M * J * Cpp, (self.ls[me], M, J), self.cs[p],
                              (M * J * Cp, J * Cp, Cp), (M * J * Cpp, J * Cpp, Cpp), tri=self.l0[me])
        n = self.L * M * J * max(_rup4(c) for c in self.cs)
        return self._run("sp_ch", n, sp_l, mk, self.L * M * J * _rup4(self.cs[me]))


def _polar_analysis_sp(x, mod, scale_kind, round_tf32=False):
    """x [B, C, K_loc, W] -> SP_l [L_loc][M][2B][Cp] (a VIEW of a peer receive buffer: consume or clone it before the
    next exchange of the same layout)  (dft, lat->chan, Legendre analysis, chan->l)."""
    B, Cn = x.shape[0], x.shape[1]
    ex = _PolarExchange.of(mod, B, Cn, x.device)
    plan = mod._peer_plan(x.device)
    x = x.contiguous()
    if ex.dft_fused(mod) and x.data_ptr() % 16 == 0:
        xm_ch = ex.dft_to_chan(x, mod._scale(x.device, scale_kind))
    else:
        xm_loc = ops.dft_fwd(x, mod.mmax, mod._scale(x.device, scale_kind))
        xm_ch = ex.lat_to_chan(xm_loc)
    sp_ch = ops.legendre_analysis(plan, xm_ch, B, ex.cs[ex.me], round_tf32=round_tf32)
    return ex.chan_to_l(sp_ch), ex


def _polar_synthesis_sp(sp_l, B, Cn, mod, scale_kind, dtype):
    """SP_l [L_loc][M][2B][Cp] -> x [B, C, K_loc, W]  (l->chan, Legendre synthesis, chan->lat, idft)."""
    ex = _PolarExchange.of(mod, B, Cn, sp_l.device)
    plan = mod._peer_plan(sp_l.device)
    sp_ch = ex.l_to_chan(sp_l)
    if ex.dft_fused(mod):
        return ex.chan_to_field(plan, sp_ch, mod.nlon, mod._scale(sp_l.device, scale_kind), dtype)
    xm_ch = ops.legendre_synthesis(plan, sp_ch, B, ex.cs[ex.me])
    xm_loc = ex.chan_to_lat(xm_ch)
    return ops.dft_inv(xm_loc, B, Cn, ex.lat[ex.me], mod.nlon, mod.mmax, mod._scale(sp_l.device, scale_kind), dtype)


def _polar_analysis(x, mod, scale_kind):
    """x [B, C, K_loc, W] -> coefficients complex64 [B, C, L_loc, M]."""
    sp_l, ex = _polar_analysis_sp(x, mod, scale_kind)
    return ops.sp_to_c64(sp_l, x.shape[0], x.shape[1], ex.ls[ex.me], mod.mmax, ex.l0[ex.me], 0)


def _polar_synthesis(c, mod, scale_kind, dtype):
    """coefficients complex64 [B, C, L_loc, M] -> x [B, C, K_loc, W]."""
    B, Cn = c.shape[0], c.shape[1]
    ex = _PolarExchange.of(mod, B, Cn, c.device)
    sp_l = ops.c64_to_sp(c.to(torch.complex64), ex.l0[ex.me], 0)
    return _polar_synthesis_sp(sp_l, B, Cn, mod, scale_kind, dtype)


class PolarSpectralConvFunction(torch.autograd.Function):
    """(x_loc, weight shard, module) -> (y_loc, residual_loc or None): the h-sharded SpectralConv (dhconv) of
    makani/models/common/spectral_convolution.py:78-82,104-107,124-148 on the peer-memory path, staying in the internal
    layouts end to end:

        dft -> [lat->chan] -> Legendre analysis -> [chan->l] -> dhconv on SP_l with l_offset (weights sharded on l, no
        communication) -> [l->chan] -> Legendre synthesis -> [chan->lat] -> inverse dft

    bf16 / fp16 / fp32 fields in and out, no complex64 tensor, no `.float()` copy, packed weight shard from the
    module's cache.  Backward is the adjoint chain with the same four exchanges in the reverse roles."""

    @staticmethod
    def forward(ctx, x, weight, mod, wp):
        fwd, inv = mod.forward_transform, mod.inverse_transform
        B, Cin, Cout = x.shape[0], mod.in_channels, mod.out_channels
        M = mod.modes_lon
        sp_l, ex = _polar_analysis_sp(x, fwd, "fwd", round_tf32=True)
        Lloc, l0 = ex.ls[ex.me], ex.l0[ex.me]
        need_grad = weight.requires_grad or x.requires_grad
        sp_l = sp_l.clone() if (need_grad or mod.scale_residual) else sp_l      # leaves the receive buffer
        residual = None
        if mod.scale_residual:
            residual = _polar_synthesis_sp(sp_l, B, Cin, inv, "one", x.dtype)
        sp2_l = ops.dhconv_fwd(sp_l, wp, B, Cin, Cout, Lloc, M, l0, 0, round_tf32=True)
        y = _polar_synthesis_sp(sp2_l, B, Cout, inv, "one", x.dtype)
        ctx.mod, ctx.in_dtype, ctx.dims = mod, x.dtype, (B, Cin, Cout, Lloc, M, l0)
        ctx.has_residual = residual is not None
        ctx.save_for_backward(sp_l, wp)
        return (y, residual) if residual is not None else (y, None)

    @staticmethod
    def backward(ctx, gy, gres):
        mod = ctx.mod
        fwd, inv = mod.forward_transform, mod.inverse_transform
        B, Cin, Cout, Lloc, M, l0 = ctx.dims
        sp_l, wp = ctx.saved_tensors
        gsp2_l, _ = _polar_analysis_sp(gy, inv, "irfft_adj", round_tf32=True)
        gw = None
        if ctx.needs_input_grad[1]:
            gw = ops.unpack_wgrad(ops.dhconv_bwd_w(sp_l, gsp2_l, B, Cin, Cout, Lloc, M, l0, 0), Cin, Cout, Lloc)
        gx = None
        if ctx.needs_input_grad[0]:
            gsp_l = ops.dhconv_bwd_x(gsp2_l, wp, B, Cin, Cout, Lloc, M, l0, 0, round_tf32=True)
            if ctx.has_residual and gres is not None:
                gr_l, _ = _polar_analysis_sp(gres, inv, "irfft_adj", round_tf32=True)
                ops.sp_band_add_(gsp_l, gr_l, B, Cin, Lloc, M, l0)
            gx = _polar_synthesis_sp(gsp_l, B, Cin, fwd, "rfft_adj", ctx.in_dtype)
        return gx, gw, None, None


def polar_spectral_conv_ok(mod, x) -> bool:
    """The fused h-sharded path applies: both transforms distributed, h-sharded over peer memory, dhconv, no bias."""
    f, i = mod.forward_transform, mod.inverse_transform
    return (isinstance(f, DistributedRealSHT) and isinstance(i, DistributedInverseRealSHT)
            and mod.operator_type == "dhconv" and f.lmax == i.lmax and f.mmax == i.mmax
            and f._peer_path(x) and i._peer_path(x))


class _PolarSHT(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, mod):
        ctx.mod, ctx.in_dtype = mod, x.dtype
        return _polar_analysis(x, mod, "fwd")

    @staticmethod
    def backward(ctx, g):
        return _polar_synthesis(g, ctx.mod, "rfft_adj", ctx.in_dtype), None


class _PolarISHT(torch.autograd.Function):
    @staticmethod
    def forward(ctx, c, mod):
        ctx.mod = mod
        return _polar_synthesis(c, mod, "one", torch.float32)

    @staticmethod
    def backward(ctx, g):
        return _polar_analysis(g, ctx.mod, "irfft_adj"), None


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

    def _peer_path(self, x: torch.Tensor) -> bool:
        """h-sharded only, CUDA + NCCL on one node: the exchanges run over peer memory (p2p.py)."""
        return (x.is_cuda and self.comm_size_azimuth == 1 and 1 < self.comm_size_polar <= 8
                and p2p.available(polar_group(), x.device))

    def _peer_plan(self, device: torch.device) -> ops.Plan:
        """Legendre table over the padded latitude order of the peer-memory path: slab p of the latitudes starts at
        column sum(round_up(lat_shapes[:p], 8)); pad columns are zero."""
        idx = device.index if device.index is not None else torch.cuda.current_device()
        plans = self.__dict__.setdefault("_peer_plans", {})
        p = plans.get(idx)
        if p is None:
            tab = self._table_slice()[1]              # w == 1 on this path: the m-slab is the whole table
            kp = sum(_rup8(k) for k in self.lat_shapes)
            padded = np.zeros(tab.shape[:2] + (kp,), dtype=np.float64)
            src = dst = 0
            for k in self.lat_shapes:
                padded[:, :, dst:dst + k] = tab[:, :, src:src + k]
                src, dst = src + k, dst + _rup8(k)
            p = ops.Plan(kp, self.nlon, self.lmax, tab.shape[0], padded, m_offset=0, device=idx)
            plans[idx] = p
        return p

    def _table_slice(self):
        m0 = sum(self.m_shapes[: self.comm_rank_azimuth])
        m1 = m0 + self.m_shapes[self.comm_rank_azimuth]
        if self.__dict__.get("_slab_registered", False):       # the registered buffer already is this rank's m-slab
            return m0, self._host_table64()
        return m0, np.ascontiguousarray(self._host_table64()[m0:m1])

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
        self._slab_registered = True

    @annotate("DistributedRealSHT.forward")
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() != 4:
            raise ValueError(f"Expected a 4-d tensor [B, C, H_local, W_local], got {x.dim()} dims")
        if self._peer_path(x):
            ops.dtype_code(x.dtype)
            return _PolarSHT.apply(x, self)
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
        self._slab_registered = True

    @annotate("DistributedInverseRealSHT.forward")
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() != 4:
            raise ValueError(f"Expected a 4-d tensor [B, C, L_local, M_local], got {x.dim()} dims")
        if self._peer_path(x):
            return _PolarISHT.apply(x, self)
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
