"""SpectralConv: drop-in for ``makani.models.common.spectral_convolution.SpectralConv``
(makani/models/common/spectral_convolution.py:43-148), with the alias ``SpectralConvS2``.

Same constructor signature, same parameter (``weight`` complex64 ``[Cin, Cout, L]`` for dhconv /
``[Cin, Cout, L, M]`` for diagonal, with the ``is_shared_mp`` / ``sharded_dims_mp`` annotations the
reference's checkpointing reads), optional ``bias``, same ``forward(x) -> (y, residual)``.

Unlike the reference, the whole chain  DFT -> Legendre -> channel contraction -> Legendre -> inverse DFT  runs
on spectral-internal device layouts without materialising a complex64 ``[B,C,L,M]`` tensor between the
stages, and its backward is a single hand-written adjoint chain (one autograd node).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from ._nvtx import annotate
from . import ops
from .sht import InverseRealSHT, RealSHT, _TransformBase

__all__ = ["SpectralConv", "SpectralConvS2"]


def _distributed_classes():
    from .distributed import DistributedInverseRealSHT, DistributedRealSHT
    return DistributedRealSHT, DistributedInverseRealSHT


class _SpectralConvFunction(torch.autograd.Function):
    """(x, weight, bias_or_None) -> (y, residual_or_None); single-device transforms only.

    forward :  x -[DFT + Legendre analysis]-> SP -[channel contraction]-> SP' -[Legendre synthesis + inverse DFT]-> y
    The two bracketed pairs run window by window over the channels (ops.sht_to_sp / ops.sp_to_field), so the Fourier
    coefficients between the DFT and the Legendre contraction never leave L2.  A constant bias is added in fp32 as the
    m = 0 Fourier coefficient, a position bias by one pass over y (spectral_convolution.py:143-146)."""

    @staticmethod
    def forward(ctx, x, weight, bias, mod, wp):
        # wp: packed dhconv weight, made by the module BEFORE entering the Function (grad mode is off in here, and
        # whether the weight is being trained decides if the packed copy may come from the cache)
        fwd, inv = mod.forward_transform, mod.inverse_transform
        dev = x.device
        B, Cin = x.shape[0], x.shape[1]
        Cout, L, M = mod.out_channels, mod.modes_lat, mod.modes_lon
        pf, pi = fwd._plan(dev), inv._plan(dev)
        one = inv._scale(dev, "one")

        sp = ops.sht_to_sp(pf, x, fwd._scale(dev, "fwd"), round_tf32=True)
        residual = None
        if mod.scale_residual:
            residual = ops.sp_to_field(pi, sp, B, Cin, inv.nlon, one, x.dtype)
        if mod.operator_type == "dhconv":
            sp2 = ops.dhconv_fwd(sp, wp, B, Cin, Cout, L, M, round_tf32=True)
        else:
            wp = None
            sp2 = ops.diag_fwd(sp, weight, B, Cin, Cout, L, M)
        dc = None
        if bias is not None and bias.numel() == Cout:
            dc = bias.detach().reshape(Cout).to(torch.float32).contiguous()
        y = ops.sp_to_field(pi, sp2, B, Cout, inv.nlon, one, x.dtype, dc_bias=dc)
        del sp2
        if bias is not None and dc is None:
            ops.add_bias_(y, bias)

        ctx.mod = mod
        ctx.packed_master = bool(getattr(mod, "_packed_master", False)) and wp is not None
        ctx.in_dtype = x.dtype
        ctx.shape = (B, Cin, Cout, L, M)
        ctx.has_residual = residual is not None
        ctx.bias_shape = None if bias is None else tuple(bias.shape)
        ctx.bias_dtype = None if bias is None else bias.dtype
        if wp is not None:
            ctx.save_for_backward(sp, wp)
        else:
            ctx.save_for_backward(sp, weight)
        if residual is None:
            return y, None
        return y, residual

    @staticmethod
    def backward(ctx, gy, gres):
        mod = ctx.mod
        fwd, inv = mod.forward_transform, mod.inverse_transform
        B, Cin, Cout, L, M = ctx.shape
        sp, wsaved = ctx.saved_tensors
        dev = gy.device
        pf, pi = fwd._plan(dev), inv._plan(dev)
        adj = inv._scale(dev, "irfft_adj")

        gbias = None
        if ctx.bias_shape is not None and ctx.needs_input_grad[2]:
            dims = (0, 2, 3) if ctx.bias_shape[-1] == 1 else (0,)
            gbias = gy.sum(dim=dims, keepdim=True, dtype=torch.float32).to(ctx.bias_dtype).reshape(ctx.bias_shape)
        gsp2 = ops.sht_to_sp(pi, gy, adj, round_tf32=True)
        gw = gwp_out = None
        if mod.operator_type == "dhconv":
            if ctx.packed_master:
                # the trained parameter IS the packed tensor: its gradient is what the GEMM writes, no layout change
                if ctx.needs_input_grad[4]:
                    gwp_out = ops.dhconv_bwd_w(sp, gsp2, B, Cin, Cout, L, M).view(wsaved.shape)
            elif ctx.needs_input_grad[1]:
                ex = getattr(mod, "_grad_exchange", None)
                if ex is not None:
                    # data-parallel replicas on one NVLink node: the GEMM writes the packed gradient straight into the
                    # exchange buffer; averaging over the ranks + unpack run on the communication stream and
                    # `weight.grad` is assigned by ex.wait()  (distributed/grad_exchange.py)
                    gwp = ops.dhconv_bwd_w(sp, gsp2, B, Cin, Cout, L, M, out=ex.gwp_out())
                    ex.reduce(mod, gwp)
                else:
                    gwp = ops.dhconv_bwd_w(sp, gsp2, B, Cin, Cout, L, M)
                    gw = ops.unpack_wgrad(gwp, Cin, Cout, L)
                del gwp
            gsp = ops.dhconv_bwd_x(gsp2, wsaved, B, Cin, Cout, L, M, round_tf32=True) if ctx.needs_input_grad[0] else None
        else:
            if ctx.needs_input_grad[1]:
                gw = ops.diag_bwd_w(sp, gsp2, B, Cin, Cout, L, M)
            gsp = ops.diag_bwd_x(gsp2, wsaved, B, Cin, Cout, L, M) if ctx.needs_input_grad[0] else None
        del gsp2
        gx = None
        if ctx.needs_input_grad[0]:
            if ctx.has_residual and gres is not None:
                # gradient of the re-projected residual: accumulated into gsp by the analysis epilogue (both tensors
                # are defined on the same band of SP; no separate add pass)
                ops.sht_to_sp(pi, gres, adj, round_tf32=True, out=gsp, accumulate=True)
            gx = ops.sp_to_field(pf, gsp, B, Cin, fwd.nlon, fwd._scale(dev, "rfft_adj"), ctx.in_dtype)
        return gx, gw, gbias, None, gwp_out


class SpectralConv(nn.Module):
    """
    Spectral convolution on the two-sphere: SHT -> per-mode channel contraction -> inverse SHT.
    Mirrors the reference module (spectral_convolution.py:43-148).
    """

    def __init__(self, forward_transform, inverse_transform, in_channels, out_channels, operator_type="diagonal",
                 separable=False, bias=False, gain=1.0):
        super().__init__()

        self.forward_transform = forward_transform
        self.inverse_transform = inverse_transform

        self.in_channels = in_channels
        self.out_channels = out_channels

        self.modes_lat = self.inverse_transform.lmax
        self.modes_lon = self.inverse_transform.mmax

        self.scale_residual = (self.forward_transform.nlat != self.inverse_transform.nlat) or (
            self.forward_transform.nlon != self.inverse_transform.nlon)
        if hasattr(self.forward_transform, "grid"):
            self.scale_residual = self.scale_residual or (self.forward_transform.grid != self.inverse_transform.grid)

        self.operator_type = operator_type
        self.separable = separable

        assert self.inverse_transform.lmax == self.modes_lat
        assert self.inverse_transform.mmax == self.modes_lon

        weight_shape = [in_channels]
        if not self.separable:
            weight_shape += [out_channels]

        DistSHT, DistISHT = _distributed_classes()
        self._distributed = isinstance(self.inverse_transform, DistISHT)
        if self._distributed:
            it = self.inverse_transform
            self.modes_lat_local = it.l_shapes[it.comm_rank_polar]
            self.modes_lon_local = it.m_shapes[it.comm_rank_azimuth]
            self.nlat_local = it.lat_shapes[it.comm_rank_polar]
            self.nlon_local = it.lon_shapes[it.comm_rank_azimuth]
            self._l_offset = sum(it.l_shapes[: it.comm_rank_polar])
            self._m_offset = sum(it.m_shapes[: it.comm_rank_azimuth])
        else:
            self.modes_lat_local = self.modes_lat
            self.modes_lon_local = self.modes_lon
            self.nlat_local = self.inverse_transform.nlat
            self.nlon_local = self.inverse_transform.nlon
            self._l_offset = 0
            self._m_offset = 0
            from .fft2 import InverseRealFFT2
            if isinstance(self.inverse_transform, InverseRealFFT2):
                # spectral_transform="fft" (sfnonet.py:541-553): Fourier coefficients have no m <= l structure, the
                # contraction kernels must not skip anything
                from . import _ext
                self._l_offset = _ext.DENSE_L_OFFSET

        if self.operator_type == "diagonal":
            weight_shape += [self.modes_lat_local, self.modes_lon_local]
        elif self.operator_type == "dhconv":
            weight_shape += [self.modes_lat_local]
        else:
            raise ValueError(f"Unsupported operator type f{self.operator_type}")

        if self.separable:
            # the reference's separable contractions ("bixy,ix->boxy", contractions.py:139-154) name an output
            # index that no operand carries; torch.einsum rejects them, so there is no behaviour to reproduce.
            raise NotImplementedError("separable SpectralConv is not supported (the reference einsum is ill-formed)")

        # initialisation of spectral_convolution.py:97-102
        # The reference multiplies a length-L scale vector into randn(*weight_shape): for dhconv that is per degree l
        # (l = 0 times sqrt 2); for diagonal [Cin, Cout, L, M] it broadcasts over the LAST axis, i.e. per order m
        # (m = 0 times sqrt 2: the coefficients that are real), and only constructs when L == M.  Same here; for
        # L != M (where the reference raises) the per-m rule is kept with a length-M vector.
        nscale = self.modes_lon_local if self.operator_type == "diagonal" else self.modes_lat_local
        scale = math.sqrt(gain / in_channels) * torch.ones(nscale, dtype=torch.complex64)
        scale[0] *= math.sqrt(2.0)
        init = scale * torch.randn(*weight_shape, dtype=torch.complex64)
        self.weight = nn.Parameter(init)

        if self.operator_type == "dhconv":
            self.weight.is_shared_mp = ["matmul", "w"]
            self.weight.sharded_dims_mp = [None for _ in weight_shape]
            self.weight.sharded_dims_mp[-1] = "h"
        else:
            self.weight.is_shared_mp = ["matmul"]
            self.weight.sharded_dims_mp = [None for _ in weight_shape]
            self.weight.sharded_dims_mp[-1] = "w"
            self.weight.sharded_dims_mp[-2] = "h"

        if bias == "constant":
            self.bias = nn.Parameter(torch.zeros(1, self.out_channels, 1, 1))
        elif bias == "position":
            self.bias = nn.Parameter(torch.zeros(1, self.out_channels, self.nlat_local, self.nlon_local))
            self.bias.is_shared_mp = ["matmul"]
            self.bias.sharded_dims_mp = [None, None, "h", "w"]

        self._wp_cache = None     # (weight version, data_ptr, packed tensor)
        self._grad_exchange = None   # set by distributed.grad_exchange.PeerGradientAverager
        self._packed_master = False  # enable_packed_master()

    # Packed planar copy of the dhconv weight (the tensor-core GEMMs cannot read the checkpoint layout).
    # While the weight is being trained (grad mode on and weight.requires_grad) it is re-packed on EVERY forward: an
    # optimizer may update the parameter without touching its version counter (apex FusedLAMB / raw-pointer updates,
    # `p.data.add_()`, EMA restores - all selectable in the reference, makani/utils/trainer.py:458-470), and a stale
    # packed copy would silently freeze the filter.  In inference (no_grad / frozen weight) the copy is cached and
    # keyed on (version, storage, device); call `invalidate()` after changing the weight behind autograd's back there.
    def _packed_weight(self, weight: torch.Tensor) -> torch.Tensor:
        key = (weight._version, weight.data_ptr(), weight.device)
        cache = self._wp_cache
        training = torch.is_grad_enabled() and weight.requires_grad
        if (cache is not None and cache[0] == key and not training
                and not torch.cuda.is_current_stream_capturing()):
            return cache[1]
        wp = ops.pack_weight(weight)
        self._wp_cache = None if training else (key, wp)
        return wp

    def invalidate(self):
        """Drop the cached packed weight (inference only; training re-packs every forward)."""
        self._wp_cache = None

    # ---- packed master copy (opt-in) ------------------------------------------------------------------------------
    # Re-packing the weight on every forward and un-packing its gradient on every backward are two pure layout changes
    # of 283 MB each way per block (15 % of the training step of the named config).  They exist only because the
    # parameter lives in the checkpoint layout complex64 [Cin, Cout, L] while the GEMMs read planar [L][2][Cout][Cin].
    # enable_packed_master() turns it around: the trained parameter becomes `weight_packed` (fp32, the GEMMs' layout),
    # its gradient is what dhconv_bwd_w writes, and `weight` is re-derived only when a state dict is taken.  Any
    # elementwise optimizer (SGD, Adam, LAMB: complex parameters are updated as view_as_real, i.e. per real number)
    # performs the same arithmetic on the packed tensor; state dicts keep the reference's key, shape and dtype
    # (`weight`, complex64 [Cin, Cout, L]; `weight_packed` never appears in them).
    # One numerical difference: the GEMMs read the fp32 master directly, so the tensor cores truncate the weight to
    # TF32 where the re-packed copy is rounded to nearest (what cuBLAS does with fp32 operands when TF32 is allowed).
    def enable_packed_master(self):
        if self._packed_master:
            return self
        if self.operator_type != "dhconv" or not self._fused_ok():
            raise NotImplementedError("enable_packed_master: dhconv on single-device transforms only")
        ops._need_cuda(self.weight, "enable_packed_master: weight")
        Cin, Cout, L = self.weight.shape
        wp = ops.pack_weight(self.weight.detach(), round_tf32=False)
        self.weight_packed = nn.Parameter(wp.view(L, 2, -1, ops._rup(Cin, 4)))
        self.weight_packed.is_shared_mp = list(self.weight.is_shared_mp)
        self.weight_packed.sharded_dims_mp = ["h", None, None, None]
        self.weight.requires_grad_(False)
        self._packed_master = True
        self._wp_cache = None
        self._register_state_dict_hook(SpectralConv._packed_state_dict_hook)
        self._register_load_state_dict_pre_hook(self._packed_load_pre_hook)
        return self

    def sync_weight_(self):
        """Refresh `weight` (checkpoint layout) from the packed master copy."""
        if self._packed_master:
            Cin, Cout, L = self.weight.shape
            with torch.no_grad():
                self.weight.copy_(ops.unpack_wgrad(self.weight_packed.detach().reshape(-1), Cin, Cout, L))
        return self.weight

    @staticmethod
    def _packed_state_dict_hook(module, state_dict, prefix, local_metadata):
        if module._packed_master:
            state_dict.pop(prefix + "weight_packed", None)
            module.sync_weight_()          # in place: the entry `weight` already in state_dict aliases the parameter
        return state_dict

    def _packed_load_pre_hook(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        key = prefix + "weight"
        if self._packed_master and key in state_dict and prefix + "weight_packed" not in state_dict:
            w = state_dict[key].to(device=self.weight_packed.device, dtype=torch.complex64)
            state_dict[prefix + "weight_packed"] = ops.pack_weight(w, round_tf32=False).view(self.weight_packed.shape)

    def _fused_ok(self):
        return (isinstance(self.forward_transform, _TransformBase) and isinstance(self.inverse_transform, _TransformBase)
                and type(self.forward_transform) is RealSHT and type(self.inverse_transform) is InverseRealSHT
                and self.forward_transform.lmax == self.modes_lat and self.forward_transform.mmax == self.modes_lon)

    def _polar_fused_ok(self, x):
        from .distributed import polar_spectral_conv_ok
        return polar_spectral_conv_ok(self, x)

    @annotate("SpectralConv.forward")
    def forward(self, x):
        dtype = x.dtype
        ops._need_cuda(x, "SpectralConv input")
        ops.dtype_code(dtype)
        if self._fused_ok():
            if self._packed_master:
                wp = self.weight_packed
            else:
                wp = self._packed_weight(self.weight) if self.operator_type == "dhconv" else None
            y, residual = _SpectralConvFunction.apply(x.contiguous(), self.weight, getattr(self, "bias", None), self, wp)
            if residual is None:
                residual = x
        elif self._distributed and not hasattr(self, "bias") and self._polar_fused_ok(x):
            from .distributed import PolarSpectralConvFunction
            y, residual = PolarSpectralConvFunction.apply(x.contiguous(), self.weight, self,
                                                          self._packed_weight(self.weight))
            if residual is None:
                residual = x
        else:
            y, residual = self._forward_composed(x)
            if hasattr(self, "bias"):
                y = y + self.bias.to(y.dtype)
        return y, residual

    # distributed (or foreign) transforms: same arithmetic, stage by stage through the module API
    def _forward_composed(self, x):
        dtype = x.dtype
        residual = x
        xs = self.forward_transform(x.float()).contiguous()
        if self.scale_residual:
            residual = self.inverse_transform(xs).to(dtype)
        from .contraction import contract_complex
        xs = contract_complex(xs, self.weight, self.operator_type, self._l_offset, self._m_offset)
        y = self.inverse_transform(xs).to(dtype)
        return y, residual


SpectralConvS2 = SpectralConv
