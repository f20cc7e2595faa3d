"""GPU test of the block rewiring (makani_b200.fuse_block): a block with the reference's attribute
layout and forward (restated from makani/models/networks/sfnonet.py:155-267 - norm0 / act_layer0 / mlp / norm1 /
drop_path / outer_skip around a SpectralConv filter; the reference class itself is not importable on the GPU box) gives
the same output and gradients before and after fusion.  Both sides run the same SpectralConv kernels, so the tolerance
is that of the fp32 pointwise arithmetic: 2e-5 forward, 2e-4 gradients (bf16 autocast: 3e-2 / 5e-2)."""
import copy

import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def mk(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import makani_b200
    return makani_b200


class _Filter(nn.Module):                     # SpectralFilterLayer (sfnonet.py:48-116): holds the conv as .filter
    def __init__(self, conv):
        super().__init__()
        self.filter = conv

    def forward(self, x):
        return self.filter(x)


class _Block(nn.Module):
    """Attribute-for-attribute stand-in of FourierNeuralOperatorBlock with the reference forward."""

    def __init__(self, mk, C, grid_in, grid_out, inner_skip, outer_skip):
        super().__init__()
        L, M = 16, 17
        fwd = mk.RealSHT(*grid_in[:2], lmax=L, mmax=M, grid=grid_in[2])
        inv = mk.InverseRealSHT(*grid_out[:2], lmax=L, mmax=M, grid=grid_out[2])
        self.norm0 = nn.InstanceNorm2d(C, eps=1e-6, affine=True, track_running_stats=False)
        if inner_skip:
            self.inner_skip = nn.Conv2d(C, C, 1, 1, bias=False)
        self.filter = _Filter(mk.SpectralConv(fwd, inv, C, C, operator_type="dhconv"))
        self.act_layer0 = nn.GELU()
        self.norm1 = nn.InstanceNorm2d(C, eps=1e-6, affine=True, track_running_stats=False)
        if outer_skip:
            self.outer_skip = nn.Conv2d(C, C, 1, 1, bias=False)
        self.mlp = nn.Sequential(nn.Conv2d(C, 2 * C, 1), nn.GELU(), nn.Conv2d(2 * C, C, 1))
        self.drop_path = nn.Identity()

    def forward(self, x):                      # sfnonet.py:239-267
        x, residual = self.filter(x)
        x = self.norm0(x)
        if hasattr(self, "inner_skip"):
            x = x + self.inner_skip(residual)
        if hasattr(self, "act_layer0"):
            x = self.act_layer0(x)
        if hasattr(self, "mlp"):
            x = self.mlp(x)
        x = self.norm1(x)
        x = self.drop_path(x)
        if hasattr(self, "outer_skip"):
            x = x + self.outer_skip(residual)
        if hasattr(self, "act_layer1"):
            x = self.act_layer1(x)
        return x


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


LG = (16, 32, "legendre-gauss")
EQ = (33, 64, "equiangular")


@pytest.mark.parametrize("grid_in,grid_out,inner,outer", [(LG, LG, False, True), (EQ, LG, False, True),
                                                          (LG, EQ, True, True), (LG, LG, True, False)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fused_block_matches_unfused(mk, grid_in, grid_out, inner, outer, dtype):
    torch.manual_seed(0)
    C = 12
    ref = _Block(mk, C, grid_in, grid_out, inner, outer).to(DEV)
    with torch.no_grad():
        for n in (ref.norm0, ref.norm1):
            n.weight.copy_(1 + 0.3 * torch.randn(C))
            n.bias.copy_(0.3 * torch.randn(C))
    fused = copy.deepcopy(ref)
    mk.fuse_block(fused)
    assert isinstance(fused.norm0, mk.FusedInstanceNorm2d) and isinstance(fused.norm1, mk.FusedInstanceNorm2d)
    assert fused._fused_outer_skip == outer and fused._fused_inner_skip == inner
    assert isinstance(fused.act_layer0, nn.GELU if inner else nn.Identity)
    assert list(fused.state_dict()) == list(ref.state_dict())
    mk.fuse_block(fused)                                                    # idempotent

    x0 = torch.randn(2, C, grid_in[0], grid_in[1], device=DEV)
    g0 = torch.randn(2, C, grid_out[0], grid_out[1], device=DEV)
    outs = []
    for m in (ref, fused):
        x = x0.clone().requires_grad_(True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=(dtype == torch.bfloat16)):
            y = m(x)
        y.backward(g0.to(y.dtype))
        outs.append((y.detach().float(), x.grad.float(), {k: p.grad.float() if not p.grad.is_complex() else p.grad
                                                         for k, p in m.named_parameters()}))
    (y_r, gx_r, gp_r), (y_f, gx_f, gp_f) = outs
    tol_y, tol_g = (2e-5, 2e-4) if dtype == torch.float32 else (3e-2, 5e-2)
    assert _rel(y_f, y_r) < tol_y
    assert _rel(gx_f, gx_r) < tol_g
    assert set(gp_r) == set(gp_f)
    # mlp.2.bias sits in front of an instance norm: its gradient is identically zero in exact arithmetic and both sides
    # hold summation noise there - require "tiny" instead of "equal"
    real = lambda t: torch.view_as_real(t) if t.is_complex() else t          # noqa: E731
    gmax = max(float(real(b).double().norm()) for b in gp_r.values())
    for k in gp_r:
        a, b = real(gp_f[k]).double(), real(gp_r[k]).double()
        if k == "mlp.2.bias":
            assert float(a.norm()) < 1e-2 * gmax and float(b.norm()) < 1e-2 * gmax, k
            continue
        err = float((a - b).norm()) / float(b.norm())
        assert err < tol_g, (k, err)


def test_fuse_block_leaves_other_norms_alone(mk):
    blk = _Block(mk, 8, LG, LG, False, True).to(DEV)
    blk.norm0 = nn.Identity()
    blk.norm1 = nn.InstanceNorm2d(8, track_running_stats=True).to(DEV)
    mk.fuse_block(blk)
    assert isinstance(blk.norm0, nn.Identity) and isinstance(blk.norm1, nn.InstanceNorm2d)
    assert isinstance(blk.act_layer0, nn.GELU) and not blk._fused_outer_skip
    y = blk(torch.randn(1, 8, 16, 32, device=DEV))
    assert y.shape == (1, 8, 16, 32)
    with pytest.raises(TypeError):
        mk.fuse_block(nn.Linear(2, 2))
