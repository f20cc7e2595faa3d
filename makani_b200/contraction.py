"""Channel contractions on complex64 ``[B, C, L, M]`` tensors - the module-level counterpart of
makani/models/common/contractions.py:121-136 and the dispatch in factorizations.py:167-200 - running on the
same kernels the fused SpectralConv uses (tcgen05 dhconv GEMM / diagonal bandwidth kernel).

Used when the transforms around the contraction are not the fusable single-device ones (the h/w-sharded
``DistributedRealSHT`` path), where ``l_offset`` / ``m_offset`` give the global degree / order of local
index 0 so the triangular m <= l structure is placed correctly.
"""
from __future__ import annotations

import torch

from . import ops

__all__ = ["contract_complex", "get_contract_fun"]


class _ContractFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, operator_type, l_offset, m_offset):
        B, Cin, L, M = x.shape
        Cout = weight.shape[1]
        sp = ops.c64_to_sp(x, l_offset, m_offset, round_tf32=(operator_type == "dhconv"))
        if operator_type == "dhconv":
            wsaved = ops.pack_weight(weight)
            sp2 = ops.dhconv_fwd(sp, wsaved, B, Cin, Cout, L, M, l_offset, m_offset)
        else:
            wsaved = weight.detach()
            sp2 = ops.diag_fwd(sp, wsaved, B, Cin, Cout, L, M, l_offset, m_offset)
        ctx.save_for_backward(sp, wsaved)
        ctx.meta = (operator_type, B, Cin, Cout, L, M, l_offset, m_offset)
        return ops.sp_to_c64(sp2, B, Cout, L, M, l_offset, m_offset)

    @staticmethod
    def backward(ctx, g):
        operator_type, B, Cin, Cout, L, M, lo, mo = ctx.meta
        sp, wsaved = ctx.saved_tensors
        gsp2 = ops.c64_to_sp(g.to(torch.complex64), lo, mo, round_tf32=(operator_type == "dhconv"))
        gx = gw = None
        if operator_type == "dhconv":
            if ctx.needs_input_grad[1]:
                gw = ops.unpack_wgrad(ops.dhconv_bwd_w(sp, gsp2, B, Cin, Cout, L, M, lo, mo), Cin, Cout, L)
            if ctx.needs_input_grad[0]:
                gx = ops.sp_to_c64(ops.dhconv_bwd_x(gsp2, wsaved, B, Cin, Cout, L, M, lo, mo), B, Cin, L, M, lo, mo)
        else:
            if ctx.needs_input_grad[1]:
                gw = ops.diag_bwd_w(sp, gsp2, B, Cin, Cout, L, M, lo, mo)
            if ctx.needs_input_grad[0]:
                gx = ops.sp_to_c64(ops.diag_bwd_x(gsp2, wsaved, B, Cin, Cout, L, M, lo, mo), B, Cin, L, M, lo, mo)
        return gx, gw, None, None, None


def contract_complex(x: torch.Tensor, weight: torch.Tensor, operator_type: str = "dhconv", l_offset: int = 0,
                     m_offset: int = 0) -> torch.Tensor:
    """einsum("bixy,iox->boxy") (dhconv) or einsum("bixy,ioxy->boxy") (diagonal) on complex64 CUDA tensors.

    Note: like the spectral data it is meant for, ``x`` is taken to vanish for m > l (the kernels skip that
    region); entries of ``x`` there are ignored and the corresponding outputs are zero.
    """
    if operator_type not in ("dhconv", "diagonal"):
        raise ValueError(f"Unkonw operator type {operator_type}")
    return _ContractFunction.apply(x.to(torch.complex64), weight, operator_type, l_offset, m_offset)


def get_contract_fun(weight, implementation="factorized", separable=False, operator_type="diagonal", complex=True):
    """Signature-compatible with factorizations.py:212 for the dense-tensor weights this path supports."""
    if not torch.is_tensor(weight):
        raise ValueError(f"Got unexpected weight type of class {weight.__class__.__name__}")
    if separable or not complex:
        raise NotImplementedError("only non-separable complex contractions are provided")

    def handle(x, w, separable=False, operator_type=operator_type):
        return contract_complex(x, w, operator_type)

    return handle
