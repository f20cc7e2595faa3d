"""GPU parity tests of makani_b200.pointwise_mlp - the 1x1-convolution stacks around the spectral filter (SURVEY 8(f)
ranks 1 and 3) as GEMMs on the NCHW tensor + fused bias / GELU kernels - against the torch modules the reference builds
(makani/models/common/layers.py:86-214: nn.Conv2d(1x1) -> nn.GELU() -> nn.Conv2d(1x1); sfnonet.py:195-198 outer_skip).

fp32: the only differences to nn.Conv2d are summation orders -> 2e-5 relative L2 on values and all gradients.
bf16 autocast: both paths round the same operands to bf16 and accumulate in fp32; they are compared with an fp64
evaluation of the same network, each to 1.5e-2 relative L2 (bf16 has 8 mantissa bits; the hidden activations are rounded
once), and the product must not be further from it than 1.5x the torch autocast path.
"""
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL32, TOL16 = 2e-5, 1.5e-2


@pytest.fixture(scope="module")
def mk(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    import makani_b200
    return makani_b200


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


class RefMLP(nn.Module):
    """layers.py:136-214 / 86-133 structure: attribute `fwd`."""

    def __init__(self, cin, hidden, cout, out_bias, dropout_slots):
        super().__init__()
        mods = [nn.Conv2d(cin, hidden, 1, bias=True), nn.GELU()]
        if dropout_slots:
            mods.append(nn.Identity())
        mods.append(nn.Conv2d(hidden, cout, 1, bias=out_bias))
        if dropout_slots:
            mods.append(nn.Identity())
        self.fwd = nn.Sequential(*mods)
        with torch.no_grad():
            for m in self.fwd:
                if isinstance(m, nn.Conv2d) and m.bias is not None:
                    m.bias.normal_(0, 0.3)

    def forward(self, x):
        return self.fwd(x)


def _grads(mod, x, g, autocast):
    for p in mod.parameters():
        p.grad = None
    x = x.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        y = mod(x)
    y.backward(g.to(y.dtype))
    return [y.detach().clone(), x.grad.clone()] + [p.grad.clone() for p in mod.parameters()]


CASES = [  # B, Cin, hidden, Cout, H, W, out_bias, dropout_slots
    (1, 16, 32, 16, 24, 48, True, True),       # MLP of a block
    (2, 7, 24, 24, 13, 21, False, False),      # encoder: odd H*W (scalar path), batch 2, no output bias
    (1, 24, 24, 5, 16, 40, False, False),      # decoder
]


@pytest.mark.parametrize("B,Cin,hid,Cout,H,W,out_bias,slots", CASES)
@pytest.mark.parametrize("autocast", [False, True])
def test_pointwise_mlp_matches_the_conv_modules(mk, B, Cin, hid, Cout, H, W, out_bias, slots, autocast):
    torch.manual_seed(3)
    ref = RefMLP(Cin, hid, Cout, out_bias, slots).to(DEV)
    x = torch.randn(B, Cin, H, W, device=DEV)
    g = torch.randn(B, Cout, H, W, device=DEV)
    want = _grads(ref, x, g, autocast)
    gold = _grads(ref.double(), x.double(), g.double(), False) if autocast else None
    ref.float()
    assert mk.fuse_pointwise(ref) and mk.fuse_pointwise(ref)            # idempotent
    got = _grads(ref, x, g, autocast)
    assert got[0].dtype == want[0].dtype and got[0].shape == want[0].shape
    names = ["y", "dx"] + [n for n, _ in ref.named_parameters()]
    for i, n in enumerate(names):
        assert got[i].dtype == want[i].dtype, n
        if autocast:
            e_got, e_ref = _rel(got[i], gold[i]), _rel(want[i], gold[i])
            assert e_got < TOL16 and e_got < 1.5 * e_ref + 1e-3, (n, e_got, e_ref)
        else:
            e = _rel(got[i], want[i])
            assert e < TOL32, (n, e)


@pytest.mark.parametrize("bias", [False, True])
@pytest.mark.parametrize("autocast", [False, True])
def test_pointwise_linear_matches_conv2d(mk, bias, autocast):
    torch.manual_seed(4)
    conv = nn.Conv2d(12, 20, 1, bias=bias).to(DEV)
    x = torch.randn(2, 12, 9, 31, device=DEV)
    g = torch.randn(2, 20, 9, 31, device=DEV)
    want = _grads(conv, x, g, autocast)
    gold = _grads(conv.double(), x.double(), g.double(), False)
    conv.float()
    assert mk.fuse_pointwise(conv)
    got = _grads(conv, x, g, autocast)
    for i in range(len(want)):
        assert got[i].dtype == want[i].dtype
        e_got, e_ref = _rel(got[i], gold[i]), _rel(want[i], gold[i])
        assert e_got < (TOL16 if autocast else TOL32) and e_got < 1.5 * e_ref + 1e-3, (i, e_got, e_ref)


def test_fuse_pointwise_leaves_other_modules_alone(mk):
    assert not mk.fuse_pointwise(nn.Conv2d(4, 4, 3, padding=1))
    assert not mk.fuse_pointwise(nn.Sequential(nn.Conv2d(4, 4, 1)))
    tanh = RefMLP(4, 8, 4, True, False)
    tanh.fwd[1] = nn.GELU(approximate="tanh")
    assert not mk.fuse_pointwise(tanh)
    with pytest.raises(RuntimeError):                       # no CPU path
        m = RefMLP(4, 8, 4, True, False)
        mk.fuse_pointwise(m)
        m(torch.randn(1, 4, 8, 8))


def test_fused_model_step_matches_the_unfused_network(mk):
    """The restated SFNO network (tools/benchlib/sfno_model.py, reference parameter names) at a small size: fuse_model
    with the pointwise path against the same network on torch convolutions / norms, fp32, values and every gradient.
    Tolerance 5e-3: an integration check, the tight bounds are in the tests above.  Both runs go through the same
    TF32 spectral path three times, and its operand rounding amplifies differences of its inputs: swapping only the
    instance norms for the fused ones (equal to 1e-6) already moves the filter-weight gradients by 6e-4
    (tools/dbg_model_noise.py on B200)."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    from benchlib import sfno_model
    torch.manual_seed(0)
    net = sfno_model.SFNO(inp_shape=(45, 90), scale_factor=3, inp_chans=5, out_chans=5, embed_dim=16, num_layers=3).to(DEV)
    x = torch.randn(2, 5, 45, 90, device=DEV)

    def run():
        for p in net.parameters():
            p.grad = None
        y = net(x)
        (y ** 2).mean().backward()
        return [y.detach()] + [p.grad.clone() for p in net.parameters()]

    want = run()
    mk.fuse_model(net)
    assert getattr(net.encoder, "_makani_b200_pointwise", False) and getattr(net.blocks[0].mlp, "_makani_b200_pointwise", False)
    got = run()
    names = ["y"] + [n for n, _ in net.named_parameters()]
    # gradients that vanish identically (a bias in front of an instance norm: mlp.fwd.3.bias) are rounding noise in both
    # runs: differences are measured against the gradient's norm plus 1e-4 of the largest gradient norm in the network
    floor = 1e-4 * max(float(torch.view_as_real(b).norm() if b.is_complex() else b.norm()) for b in want[1:])
    for n, a, b in zip(names, got, want):
        a, b = (torch.view_as_real(a), torch.view_as_real(b)) if a.is_complex() else (a, b)
        err = float((a.double() - b.double()).norm() / (b.double().norm() + (floor if n != "y" else 0.0)))
        assert err < 5e-3, (n, err)
