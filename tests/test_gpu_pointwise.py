"""GPU parity tests of the fused instance-norm kernels (csrc/pointwise.cu, makani_b200.FusedInstanceNorm2d) against
the torch modules the reference block composes (sfnonet.py:239-267: nn.InstanceNorm2d(eps=1e-6, affine=True), nn.GELU,
the outer-skip add) evaluated in fp64 on the CPU, forward and backward.

Stated tolerances (rel-L2 over the whole tensor): fp32 I/O 2e-6 forward / 1e-5 gradients (fp32 sums over up to 1e6
elements); bf16 I/O 1e-2 (rounding of the stored result dominates)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def mk(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import makani_b200
    return makani_b200


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def _reference(x, w, b, skip, eps, act):
    """fp64 CPU composition of the reference's modules."""
    y = F.instance_norm(x, weight=w, bias=b, eps=eps)
    if act == "gelu":
        y = F.gelu(y)
    if skip is not None:
        y = y + skip
    return y


CASES = [
    # B, C, H, W, affine, act, skip
    (2, 5, 12, 24, True, "gelu", False),       # vector path (H*W multiple of 8)
    (2, 5, 12, 24, True, "none", True),        # norm1(x) + outer_skip(residual)
    (1, 3, 7, 9, True, "gelu", False),         # 63 elements per plane: scalar path
    (1, 3, 7, 9, False, "none", True),
    (3, 4, 33, 64, False, "gelu", False),      # affine=False
    (1, 16, 240, 480, True, "gelu", False),    # internal grid
    (1, 8, 721, 1440, True, "none", True),     # full-resolution grid (block 7)
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,C,H,W,affine,act,use_skip", CASES)
def test_fused_instance_norm_forward_backward(mk, B, C, H, W, affine, act, use_skip, dtype):
    torch.manual_seed(B * 1000 + C * 10 + H)
    eps = 1e-6
    mod = mk.FusedInstanceNorm2d(C, eps=eps, affine=affine, act=act).to(DEV)
    if affine:
        with torch.no_grad():
            mod.weight.copy_(1.0 + 0.5 * torch.randn(C))
            mod.bias.copy_(0.5 * torch.randn(C))
    # per-plane offsets and scales so that the statistics matter
    x0 = (torch.randn(B, C, H, W) * (0.5 + torch.rand(B, C, 1, 1)) + 3.0 * torch.randn(B, C, 1, 1)).to(dtype)
    s0 = torch.randn(B, C, H, W).to(dtype) if use_skip else None
    g0 = torch.randn(B, C, H, W).to(dtype)

    x = x0.to(DEV).requires_grad_(True)
    skip = s0.to(DEV).requires_grad_(True) if use_skip else None
    y = mod(x, skip) if use_skip else mod(x)
    assert y.dtype == dtype and y.shape == x.shape
    y.backward(g0.to(DEV))

    xr = x0.double().requires_grad_(True)
    sr = s0.double().requires_grad_(True) if use_skip else None
    wr = mod.weight.detach().double().cpu().requires_grad_(True) if affine else None
    br = mod.bias.detach().double().cpu().requires_grad_(True) if affine else None
    yr = _reference(xr, wr, br, sr, eps, act)
    yr.backward(g0.double())

    tol_f, tol_b = (2e-6, 1e-5) if dtype == torch.float32 else (1e-2, 1e-2)
    assert _rel(y.detach().cpu(), yr.detach()) < tol_f
    assert _rel(x.grad.cpu(), xr.grad) < tol_b
    if use_skip:
        assert torch.equal(skip.grad.cpu(), g0)
    if affine:
        # parameter gradients are fp32 sums of products of the ROUNDED inputs: compare at the input precision
        assert _rel(mod.weight.grad.cpu(), wr.grad) < max(tol_b, 1e-5)
        assert _rel(mod.bias.grad.cpu(), br.grad) < max(tol_b, 1e-5)


def test_fp16_io(mk):
    """fp16 in / out (third dtype of the C ABI): same kernels, tolerance 2e-3 (fp16 rounding of the stored result)."""
    torch.manual_seed(4)
    mod = mk.FusedInstanceNorm2d(6, eps=1e-6, affine=True, act="gelu").to(DEV)
    with torch.no_grad():
        mod.weight.copy_(1.0 + 0.5 * torch.randn(6))
        mod.bias.copy_(0.5 * torch.randn(6))
    x0 = (torch.randn(2, 6, 24, 48) * 1.5 + 0.7).to(torch.float16)
    g0 = torch.randn(2, 6, 24, 48).to(torch.float16)
    x = x0.to(DEV).requires_grad_(True)
    y = mod(x)
    assert y.dtype == torch.float16
    y.backward(g0.to(DEV))
    xr = x0.double().requires_grad_(True)
    yr = _reference(xr, mod.weight.detach().double().cpu(), mod.bias.detach().double().cpu(), None, 1e-6, "gelu")
    yr.backward(g0.double())
    assert _rel(y.detach().cpu(), yr.detach()) < 2e-3
    assert _rel(x.grad.cpu(), xr.grad) < 2e-3


def test_large_mean_does_not_cancel(mk):
    """fields with |mean| >> std (e.g. geopotential, temperature in K): the shifted sums keep the variance."""
    torch.manual_seed(1)
    x0 = 300.0 + 0.25 * torch.randn(1, 4, 64, 128)
    mod = mk.FusedInstanceNorm2d(4, eps=1e-6).to(DEV)
    y = mod(x0.to(DEV))
    yr = F.instance_norm(x0.double(), eps=1e-6)
    assert _rel(y.cpu(), yr) < 1e-4            # input quantisation: ulp(300) / 0.25 = 1.2e-4 per element before averaging


def test_matches_torch_modules_state_dict_and_cuda(mk):
    """same parameters as nn.InstanceNorm2d: load its state dict, compare with the torch CUDA modules in fp32."""
    torch.manual_seed(2)
    ref = torch.nn.InstanceNorm2d(6, eps=1e-6, affine=True, track_running_stats=False).to(DEV)
    with torch.no_grad():
        ref.weight.normal_()
        ref.bias.normal_()
    mod = mk.FusedInstanceNorm2d(6, eps=1e-6, affine=True, act="gelu").to(DEV)
    mod.load_state_dict(ref.state_dict(), strict=True)
    x = torch.randn(2, 6, 40, 80, device=DEV)
    assert _rel(mod(x), torch.nn.GELU()(ref(x))) < 2e-6


def test_errors(mk):
    mod = mk.FusedInstanceNorm2d(4, act="gelu").to(DEV)
    with pytest.raises(RuntimeError, match="no CPU path"):
        mod(torch.randn(1, 4, 8, 8))
    with pytest.raises(ValueError):
        mod(torch.randn(1, 4, 8, 8, device=DEV), torch.randn(1, 4, 8, 8, device=DEV))     # skip with an activation
    with pytest.raises(ValueError):
        mod(torch.randn(1, 3, 8, 8, device=DEV))
    with pytest.raises(TypeError):
        mod(torch.randn(1, 4, 8, 8, device=DEV, dtype=torch.float64))
    with pytest.raises(ValueError):
        mk.FusedInstanceNorm2d(4, act="relu")


def test_full_size_properties(mk):
    """C = 384 at both grids of the named config, bf16: every output plane of act='none' has mean beta_c and
    variance gamma_c^2 (size-independent property of the normalisation)."""
    torch.manual_seed(3)
    for H, W in ((240, 480), (721, 1440)):
        C = 384
        mod = mk.FusedInstanceNorm2d(C, eps=1e-6, affine=True).to(DEV)
        with torch.no_grad():
            mod.weight.copy_(0.5 + torch.rand(C))
            mod.bias.copy_(torch.randn(C))
        x = (torch.randn(1, C, H, W, device=DEV) * 2.0 + 5.0).to(torch.bfloat16)
        y = mod(x).float()
        var, mean = torch.var_mean(y, dim=(-2, -1), unbiased=False)
        assert float((mean[0] - mod.bias).abs().max()) < 2e-2
        assert float((var[0].sqrt() / mod.weight - 1).abs().max()) < 1e-2
        del x, y
