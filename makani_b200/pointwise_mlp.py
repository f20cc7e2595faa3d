"""1x1-convolution stacks of the SFNO network without layout changes - SURVEY 8(f) ranks 1 and 3.

Besides the spectral filter, a ``FourierNeuralOperatorBlock`` (makani/models/networks/sfnonet.py:239-267) and the
network around it (sfnonet.py:356-363, 451-467, 637-638) consist of 1x1 convolutions on NCHW fields:

    encoder / decoder  EncoderDecoder : Conv2d(1x1, bias) -> GELU -> Conv2d(1x1, no bias)      layers.py:86-133
    block.mlp          MLP            : Conv2d(1x1, bias) -> GELU -> Conv2d(1x1, bias)         layers.py:136-214
    block.outer_skip, residual_transform : Conv2d(1x1, no bias)                                sfnonet.py:195-198, 464-468

``nn.Conv2d`` sends them through cuDNN, which transposes every operand NCHW -> NHWC and back and adds the bias in a
separate unvectorised pass: in the whole-model step of the named configuration (bench.py `full_model`, 721x1440,
bf16 autocast) those transposes and bias passes alone took 18 of 64 ms (profiles/r2j_full_model_kernels.txt).  A 1x1
convolution on an NCHW tensor needs no layout change at all: with the field viewed as ``[C, H*W]`` it is the plain GEMM
``W[Cout, Cin] @ X[Cin, H*W]`` whose long dimension is contiguous.  Here

    GEMMs          torch.mm on those views (library GEMMs, fp32 accumulation; 3 per layer with the backward)
    in between     csrc/pointwise.cu: ``sfno_bias_act`` (bias + exact GELU, one pass) and ``sfno_bias_act_bwd``
                   (GELU gradient and the bias-gradient sums in the same pass)

``fuse_pointwise(module)`` reroutes ``forward`` of the reference's own module instances (parameters, their names and
the ``is_shared_mp`` annotations stay as they are, so state dicts and the gradient hooks are unaffected);
``makani_b200.fuse_model`` applies it to the encoder, the decoder, the big-skip convolution and every block.

CUDA only: no CPU path.
"""
from __future__ import annotations

import ctypes as C
import types
from typing import Optional

import torch
import torch.nn as nn

from . import _ext
from ._nvtx import annotate
from .ops import _need_cuda, _ptr, _stream, dtype_code

__all__ = ["pointwise_linear", "pointwise_mlp", "fuse_pointwise"]

_SEG = _ext.INSTNORM_SEGMENTS


def _opt(t: Optional[torch.Tensor]):
    return C.c_void_p(0) if t is None else _ptr(t)


def _bias_act_(t: torch.Tensor, bias32, act, out=None):
    """out = act(t + bias[c]) on [B, C, HW]; in place when out is None."""
    B, Cn, hw = t.shape
    out = t if out is None else out
    _ext.check(_ext.load().sfno_bias_act(_ptr(t), dtype_code(t.dtype), B * Cn, Cn, hw, _opt(bias32), act, _ptr(out),
                                         _stream()), "sfno_bias_act")
    return out


def _bias_act_bwd(g: torch.Tensor, pre, bias32, act, want_dx: bool):
    """(dx written over g when want_dx, d bias [C] fp32)"""
    B, Cn, hw = g.shape
    part = torch.empty(B * Cn, _SEG, dtype=torch.float32, device=g.device)
    _ext.check(_ext.load().sfno_bias_act_bwd(_ptr(g), _opt(pre), dtype_code(g.dtype), B * Cn, Cn, hw, _opt(bias32), act,
                                             _ptr(g) if want_dx else C.c_void_p(0), _ptr(part), _stream()),
               "sfno_bias_act_bwd")
    return part.view(B, Cn, _SEG).sum(dim=(0, 2))


# GEMMs per batch entry with torch.mm on 2-D views: every operand and result is a plain row-major matrix (or its
# transpose, which cuBLAS takes as an operand flag).  torch.matmul(2-D, 3-D) must not be used here: it folds the batch
# by transposing and hands back a transposed view, and the copy that makes it contiguous again costs more than the GEMM.
def _apply_w(w: torch.Tensor, x3: torch.Tensor) -> torch.Tensor:
    """y[b] = w @ x[b] : [B, Cout, HW]   (w may be a transposed view)"""
    y = torch.empty(x3.shape[0], w.shape[0], x3.shape[2], dtype=x3.dtype, device=x3.device)
    for b in range(x3.shape[0]):
        torch.mm(w, x3[b], out=y[b])
    return y


def _wgrad(g3: torch.Tensor, x3: torch.Tensor) -> torch.Tensor:
    """sum_b g[b] @ x[b]^T : [Cout, Cin]  (contraction over the pixels)"""
    gw = torch.mm(g3[0], x3[0].t())
    for b in range(1, g3.shape[0]):
        gw.addmm_(g3[b], x3[b].t())
    return gw


class _PointwiseLinear(torch.autograd.Function):
    """y = W x (+ b) per pixel; x [B, Cin, HW] contiguous, W [Cout, Cin], b fp32 [Cout] or None."""

    @staticmethod
    def forward(ctx, x3, w, b32):
        y = _apply_w(w, x3)
        if b32 is not None:
            _bias_act_(y, b32, _ext.ACT_NONE)
        ctx.save_for_backward(x3, w)
        ctx.has_bias = b32 is not None
        return y

    @staticmethod
    def backward(ctx, g):
        x3, w = ctx.saved_tensors
        g = g.contiguous()
        gx = gw = gb = None
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = _bias_act_bwd(g, None, None, _ext.ACT_NONE, want_dx=False)
        if ctx.needs_input_grad[1]:
            gw = _wgrad(g, x3)
        if ctx.needs_input_grad[0]:
            gx = _apply_w(w.t(), g)
        return gx, gw, gb


class _PointwiseMLP(torch.autograd.Function):
    """y = W2 gelu(W1 x + b1) (+ b2) per pixel."""

    @staticmethod
    def forward(ctx, x3, w1, b1, w2, b2):
        pre = _apply_w(w1, x3)                                      # [B, hidden, HW]
        act = _bias_act_(pre, b1, _ext.ACT_GELU, out=torch.empty_like(pre))
        y = _apply_w(w2, act)
        if b2 is not None:
            _bias_act_(y, b2, _ext.ACT_NONE)
        ctx.save_for_backward(x3, w1, w2, pre, act, *([b1] if b1 is not None else []))
        ctx.has_b = (b1 is not None, b2 is not None)
        return y

    @staticmethod
    def backward(ctx, g):
        saved = ctx.saved_tensors
        x3, w1, w2, pre, act = saved[:5]
        b1 = saved[5] if ctx.has_b[0] else None
        g = g.contiguous()
        gx = gw1 = gb1 = gw2 = gb2 = None
        if ctx.has_b[1] and ctx.needs_input_grad[4]:
            gb2 = _bias_act_bwd(g, None, None, _ext.ACT_NONE, want_dx=False)
        if ctx.needs_input_grad[3]:
            gw2 = _wgrad(g, act)
        d = _apply_w(w2.t(), g)                                     # d act, then d pre in place
        s = _bias_act_bwd(d, pre, b1, _ext.ACT_GELU, want_dx=True)
        if ctx.has_b[0] and ctx.needs_input_grad[2]:
            gb1 = s
        if ctx.needs_input_grad[1]:
            gw1 = _wgrad(d, x3)
        if ctx.needs_input_grad[0]:
            gx = _apply_w(w1.t(), d)
        return gx, gw1, gb1, gw2, gb2


def _prepare(x: torch.Tensor):
    _need_cuda(x, "x")
    if x.dim() != 4:
        raise ValueError(f"expected [B, C, H, W], got {tuple(x.shape)}")
    dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
    if dt not in (torch.float32, torch.bfloat16, torch.float16):
        raise TypeError(f"makani_b200: unsupported dtype {dt}")
    B, Cn, H, W = x.shape
    x3 = x.to(dt).contiguous().view(B, Cn, H * W)
    # 16-bit GEMM operands want leading dimensions that are multiples of 16 bytes: with 73 input channels (the ERA5
    # variables: encoder, big skip) the library falls back to kernels three times slower.  Zero channels are appended
    # to x and zero columns to the weight (one extra pass over the 73-channel field, 2 % of the encoder's traffic).
    if dt != torch.float32 and Cn % 8:
        x3 = torch.nn.functional.pad(x3, (0, 0, 0, 8 - Cn % 8))
    return x3, dt


def _w2d(w: torch.Tensor, dt, cin: int) -> torch.Tensor:
    """2-D view of a 1x1 convolution weight in the compute dtype; cin = channels of the (possibly padded) input."""
    if w.dim() == 4:
        if w.shape[2:] != (1, 1):
            raise ValueError(f"only 1x1 convolution weights, got {tuple(w.shape)}")
        w = w.view(w.shape[0], w.shape[1])
    w = w.to(dt)
    if dt != torch.float32 and w.shape[1] % 8 and w.shape[1] + 8 - w.shape[1] % 8 == cin:
        w = torch.nn.functional.pad(w, (0, cin - w.shape[1]))
    if w.shape[1] != cin:
        raise ValueError(f"weight {tuple(w.shape)} does not match {cin} input channels")
    return w


def _b32(b):
    return None if b is None else b.to(torch.float32)


@annotate("pointwise_linear")
def pointwise_linear(x, weight, bias=None):
    """``F.conv2d(x, weight, bias)`` for a 1x1 kernel (weight [Cout, Cin, 1, 1] or [Cout, Cin]) on a CUDA tensor
    [B, Cin, H, W]; under autocast the GEMM runs in the autocast dtype like the convolution would."""
    x3, dt = _prepare(x)
    y = _PointwiseLinear.apply(x3, _w2d(weight, dt, x3.shape[1]), _b32(bias))
    return y.view(x.shape[0], -1, x.shape[2], x.shape[3])


@annotate("pointwise_mlp")
def pointwise_mlp(x, weight1, bias1, weight2, bias2=None):
    """``conv1x1(gelu(conv1x1(x, weight1, bias1)), weight2, bias2)`` with the exact (erf) GELU of ``nn.GELU()``."""
    x3, dt = _prepare(x)
    w1 = _w2d(weight1, dt, x3.shape[1])
    w2 = _w2d(weight2, dt, w1.shape[0])
    y = _PointwiseMLP.apply(x3, w1, _b32(bias1), w2, _b32(bias2))
    return y.view(x.shape[0], -1, x.shape[2], x.shape[3])


# ------------------------------------------------------------------------------------------------ rewiring
def _is_conv1x1(m) -> bool:
    return (isinstance(m, nn.Conv2d) and m.kernel_size == (1, 1) and m.stride == (1, 1) and m.padding == (0, 0)
            and m.dilation == (1, 1) and m.groups == 1 and m.padding_mode == "zeros")


def _exact_gelu(m) -> bool:
    return isinstance(m, nn.GELU) and getattr(m, "approximate", "none") == "none"


def _mlp_pattern(seq):
    """(conv1, conv2) if `seq` is Conv1x1(bias) -> GELU -> [Identity]* -> Conv1x1 -> [Identity]*, else None."""
    if not isinstance(seq, nn.Sequential):
        return None
    mods = [m for m in seq if not isinstance(m, nn.Identity)]
    if len(mods) == 3 and _is_conv1x1(mods[0]) and _exact_gelu(mods[1]) and _is_conv1x1(mods[2]):
        return mods[0], mods[2]
    return None


def _linear_forward(self, x):
    return pointwise_linear(x, self.weight, self.bias)


def _mlp_forward(self, x):
    c1, c2 = self._pointwise_convs
    return pointwise_mlp(x, c1.weight, c1.bias, c2.weight, c2.bias)


def fuse_pointwise(module: nn.Module) -> bool:
    """Reroute ``module.forward`` in place if it is a 1x1 ``nn.Conv2d`` or a reference ``MLP`` / ``EncoderDecoder``
    (attribute ``fwd``: Conv1x1 + bias -> exact GELU -> Conv1x1, dropouts being ``nn.Identity``; gradient
    checkpointing off).  Returns whether it was rewired; anything else is left alone.  Idempotent."""
    if getattr(module, "_makani_b200_pointwise", False):
        return True
    if _is_conv1x1(module):
        module.forward = types.MethodType(_linear_forward, module)
    else:
        convs = _mlp_pattern(getattr(module, "fwd", None))
        if convs is None or getattr(module, "checkpointing", 0) >= 2:
            return False
        module._pointwise_convs = convs
        module.forward = types.MethodType(_mlp_forward, module)
    module._makani_b200_pointwise = True
    return True
