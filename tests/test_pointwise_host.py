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


def _sharded_worker(rank, world, port):
    import os
    import torch.distributed as dist
    import torch.nn.functional as F
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import fake_pointwise_lib
        fake = fake_pointwise_lib.install()
        import makani_b200 as mk
        torch.manual_seed(11)                                          # same full tensors on every rank
        B, Cn, K, W = 2, 3, 37, 10
        x_full = torch.randn(B, Cn, K, W) * torch.linspace(0.5, 2.0, K).reshape(1, 1, K, 1) + 1.5
        g_full = torch.randn(B, Cn, K, W)
        lat = [19, 18]                                                 # uneven slabs: compute_split_shapes(37, 2)
        for act in ("gelu", "none"):
            mod = mk.DistributedFusedInstanceNorm2d(Cn, eps=1e-6, affine=True, act=act, group=dist.group.WORLD)
            with torch.no_grad():
                mod.weight.copy_(torch.tensor([1.5, 0.5, -1.0]))
                mod.bias.copy_(torch.tensor([0.1, -0.2, 0.3]))
            x = torch.split(x_full, lat, dim=2)[rank].contiguous().requires_grad_(True)
            y = mod(x)
            y.backward(torch.split(g_full, lat, dim=2)[rank].contiguous())

            xr = x_full.double().requires_grad_(True)
            wr = mod.weight.detach().double().requires_grad_(True)
            br = mod.bias.detach().double().requires_grad_(True)
            yr = F.instance_norm(xr, weight=wr, bias=br, eps=1e-6)
            yr = F.gelu(yr) if act == "gelu" else yr
            yr.backward(g_full.double())
            sl = lambda t: torch.split(t, lat, dim=2)[rank]             # noqa: E731
            rel = lambda a, b: float((a.double() - b).norm() / b.norm())   # noqa: E731
            assert rel(y.detach(), sl(yr.detach())) < 1e-5, act
            assert rel(x.grad, sl(xr.grad)) < 1e-5, act
            # parameter gradients are this rank's partial sums (is_shared_mp = ["spatial"]: the hooks add them up)
            gw, gb = mod.weight.grad.clone(), mod.bias.grad.clone()
            dist.all_reduce(gw)
            dist.all_reduce(gb)
            assert rel(gw, wr.grad) < 1e-5 and rel(gb, br.grad) < 1e-5, act
        assert fake.calls.count("stats") == 2 and fake.calls.count("bwd_apply") == 2
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_sharded_instance_norm_host_logic_gloo():
    """World-size-2 gloo run of DistributedFusedInstanceNorm2d's host logic (moment merge over uneven latitude slabs,
    backward all-reduce of the two per-plane sums, partial parameter gradients) with the four C entry points emulated on
    the CPU (tests/fake_pointwise_lib.py), against F.instance_norm (+ GELU) on the full field."""
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_sharded_worker, args=(2, port), nprocs=2, join=True)
