"""CPU tests of the host logic around the fused instance norm: moment merging (the Welford / Chan update the reference
applies in makani/mpu/layer_norm.py:66-80), module surface and state-dict compatibility with nn.InstanceNorm2d."""
import pytest
import torch


def test_merge_moments_matches_direct_statistics():
    from makani_b200.instance_norm import merge_moments
    torch.manual_seed(0)
    x = torch.randn(6, 1000, dtype=torch.float64) * torch.linspace(0.5, 4, 6)[:, None] + torch.arange(6.0)[:, None] * 10
    sizes = [181 + 181, 181 + 178, 279]                      # uneven slabs
    parts = torch.split(x, sizes, dim=1)
    mean = torch.stack([p.mean(dim=1) for p in parts], dim=-1)
    var = torch.stack([p.var(dim=1, unbiased=False) for p in parts], dim=-1)
    count = torch.tensor([float(s) for s in sizes], dtype=torch.float64)
    m, v, n = merge_moments(mean, var, count)
    torch.testing.assert_close(m, x.mean(dim=1), rtol=1e-12, atol=1e-12)
    torch.testing.assert_close(v, x.var(dim=1, unbiased=False), rtol=1e-12, atol=1e-12)
    assert torch.all(n == 1000)
    # one rank: identity
    m1, v1, n1 = merge_moments(mean[..., :1], var[..., :1], count[:1])
    torch.testing.assert_close(m1, mean[..., 0])
    torch.testing.assert_close(v1, var[..., 0])


def test_module_surface_and_state_dict():
    import makani_b200 as mk
    ref = torch.nn.InstanceNorm2d(8, eps=1e-6, affine=True, track_running_stats=False)
    mod = mk.FusedInstanceNorm2d(8, eps=1e-6, affine=True, track_running_stats=False, act="gelu")
    assert {k: tuple(v.shape) for k, v in mod.state_dict().items()} == \
           {k: tuple(v.shape) for k, v in ref.state_dict().items()}
    mod.load_state_dict(ref.state_dict(), strict=True)
    plain = mk.FusedInstanceNorm2d(8)
    assert plain.weight is None and plain.bias is None and len(plain.state_dict()) == 0
    d = mk.DistributedFusedInstanceNorm2d(8, eps=1e-6, affine=True)
    assert d.weight.is_shared_mp == ["spatial"] and d.bias.is_shared_mp == ["spatial"]
    assert d._group() is None                                # no process groups: behaves like the local module
    with pytest.raises(NotImplementedError):
        mk.FusedInstanceNorm2d(8, track_running_stats=True)
    with pytest.raises(RuntimeError, match="no CPU path"):
        mod(torch.randn(1, 8, 4, 4))


def test_fast_erf_host_accuracy(tmp_path):
    """pointwise_math.cuh compiled for the host: the A&S 7.1.26 erf (the -DSFNO_FAST_ERF=1 variant of the GELU kernels)
    and the default path against double-precision erf on a dense grid (bounds in tests/cpp/gelu_host_test.cpp)."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = os.path.join(root, "tests", "cpp", "gelu_host_test.cpp")
    for flag in ("-DSFNO_FAST_ERF=1", "-DSFNO_FAST_ERF=0"):
        exe = str(tmp_path / ("gelu" + flag[-1]))
        subprocess.run(["g++", "-O2", "-std=c++17", flag, src, "-o", exe], check=True)
        r = subprocess.run([exe], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout
