"""CPU tests of bench.py's host logic (no kernels): the chained step's autograd plumbing with stand-in blocks that
have the SpectralConv call signature, the algorithmic-byte formulas of DESIGN.md, and the JSON contract of the
`--impl reference` arm's line builder."""
import importlib.util
import os

import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class _StandInBlock(nn.Module):
    """(y, residual) = (resize(x) * w, resize(x) or x) - the call signature of SpectralConv.forward
    (spectral_convolution.py:124-148), cheap enough for the CPU."""

    def __init__(self, n_in, n_out):
        super().__init__()
        self.weight = nn.Parameter(torch.randn(3, dtype=torch.complex64))
        self.n_in, self.n_out = n_in, n_out

    def _resize(self, x):
        if self.n_in == self.n_out:
            return x
        return torch.nn.functional.interpolate(x, size=(self.n_out, 2 * self.n_out), mode="bilinear")

    def forward(self, x):
        assert x.shape[-2] == self.n_in
        xr = self._resize(x)
        y = xr * self.weight.abs().sum()
        return y, (xr if self.n_in != self.n_out else x)


def test_chained_step_matches_manual_composition():
    b = _bench()
    torch.manual_seed(0)
    big, small = 12, 4
    sizes = [(big, small)] + [(small, small)] * 6 + [(small, big)]
    blocks = [_StandInBlock(i, o) for i, o in sizes]
    x0 = torch.rand(1, 3, big, 2 * big, requires_grad=True)
    grads = []
    for k, (i, o) in enumerate(sizes):
        if i != o:
            g = torch.rand(1, 3, o, 2 * o) - 0.5
            grads.append((g, g.clone()))
        else:
            grads.append((None, None))
    chk = b.gpu_step_chained(blocks, x0, grads)
    assert chk.shape == (2,) and torch.isfinite(chk).all()
    got_x = x0.grad.clone()
    got_w = [blk.weight.grad.clone() for blk in blocks]

    # manual composition: loss = <y7, gy7> + <res7, gr7> + <res0, gr0>
    x0.grad = None
    for blk in blocks:
        blk.weight.grad = None
    x = x0
    loss = 0.0
    for blk, (gy, gr) in zip(blocks, grads):
        y, res = blk(x)
        if gr is not None:
            loss = loss + (res * gr).sum()
        x = y
    loss = loss + (x * grads[-1][0]).sum()
    loss.backward()
    torch.testing.assert_close(got_x, x0.grad)
    for a, blk in zip(got_w, blocks):
        torch.testing.assert_close(a, blk.weight.grad)

    # a second step starts from clean gradients (no accumulation across steps)
    b.gpu_step_chained(blocks, x0, grads)
    torch.testing.assert_close(x0.grad, got_x)


def test_make_grads_and_inputs_shapes(monkeypatch):
    b = _bench()
    b = b.W                                                 # tools/benchlib/workload.py
    monkeypatch.setattr(b, "C_EMBED", 2)
    monkeypatch.setattr(b, "GRID_IN", (9, 16, "equiangular"))
    monkeypatch.setattr(b, "GRID_INT", (4, 8, "legendre-gauss"))
    cpu = torch.device("cpu")
    ind = b.make_grads(1, torch.float32, cpu)
    assert len(ind) == 8 and all(g is not None for g, _ in ind)
    assert ind[0][1] is not None and ind[7][1] is not None and all(r is None for _, r in ind[1:7])
    ch = b.make_grads(1, torch.float32, cpu, chained=True)
    assert [g is not None for g, _ in ch] == [True] + [False] * 6 + [True]
    assert ch[7][0].shape == (1, 2, 9, 16) and ch[0][1].shape == (1, 2, 4, 8)
    assert len(b.make_inputs(1, torch.float32, False, chained=True)) == 1
    xs = b.make_inputs(1, torch.float32, False)
    assert [tuple(x.shape[-2:]) for x in xs] == [(9, 16)] + [(4, 8)] * 7


def test_algorithmic_bytes_formulas():
    """DESIGN.md section 4 / SURVEY 8(d): bytes per launch at the named config."""
    b = _bench()
    tri = sum(min(l + 1, 241) for l in range(240))
    assert tri == 240 * 241 // 2
    kb = b.W.kernel_bytes
    a = dict(B=1, C=384, K=240, W=480, M=241, esize=2)
    assert kb("dft_fwd", a) == 384 * 240 * (480 * 2 + 241 * 8)
    leg = kb("legendre_analysis", dict(B=1, C=384, K=240, L=240, M=241))
    assert leg == 384 * 8 * (240 * 241 + tri) + 4 * 241 * 240 * 240
    assert abs(leg / 1e6 - 322) < 1.0                                   # DESIGN: 322 MB
    # channel windows: the pieces add up to the whole, the table is charged once
    parts = sum(kb("legendre_analysis", dict(B=1, C=128, K=240, L=240, M=241, c0=c0)) for c0 in (0, 128, 256))
    assert parts == leg
    dh = kb("dhconv_fwd", dict(B=1, Cin=384, Cout=384, L=240, M=241))
    assert abs(dh / 1e6 - 461) < 1.0                                    # DESIGN: 461 MB
    assert kb("pack_weight", dict(Cin=384, Cout=384, L=240)) == 16 * 384 * 384 * 240
    # one fused internal block forward, fp32 I/O: 353.9 b + 394.2 MB; the bf16-I/O variant; the whole step
    g = b.W.GRID_INT
    assert abs(b.W.block_forward_bytes(g, g, 1, esize=4) / 1e6 - 748.1) < 0.1
    assert abs(b.W.block_forward_bytes(g, g, 8, esize=4) / 1e6 - 3225.3) < 0.1
    assert abs(b.W.block_forward_bytes(g, g, 1, esize=2) / 1e6 - 571.1) < 0.1
    assert abs(b.W.step_bytes() / 1e9 - 17.87) < 0.01


def _chained_dp_worker(rank, world, port):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from makani_b200.distributed.grad_sync import GradientAverager
        b = _bench()
        torch.manual_seed(0)                                   # same weights on every replica
        big, small = 12, 4
        sizes = [(big, small)] + [(small, small)] * 6 + [(small, big)]
        blocks = [_StandInBlock(i, o) for i, o in sizes]
        ref = [torch.nn.Parameter(blk.weight.detach().clone()) for blk in blocks]
        sync = GradientAverager([blk.weight for blk in blocks])
        gen = torch.Generator().manual_seed(10 + rank)         # different data per replica
        grads = []
        for i, o in sizes:
            if i != o:
                g = torch.rand(1, 3, o, 2 * o, generator=gen) - 0.5
                grads.append((g, g.clone()))
            else:
                grads.append((None, None))
        for _ in range(2):                                      # two steps: hooks stay armed
            x0 = torch.rand(1, 3, big, 2 * big, generator=gen).requires_grad_(True)
            b.gpu_step_chained(blocks, x0, grads)               # the all-reduces start from the hooks during backward
            sync.wait()
            # the replica's own gradient, without hooks, then averaged by hand
            for blk, r in zip(blocks, ref):
                blk.weight, keep = r, blk.weight
                blk._keep = keep
            for r in ref:
                r.grad = None
            b.gpu_step_chained(blocks, x0.detach().clone().requires_grad_(True), grads)
            for blk in blocks:
                blk.weight = blk._keep
            for blk, r in zip(blocks, ref):
                mine = torch.view_as_real(r.grad).contiguous()
                parts = [torch.zeros_like(mine) for _ in range(world)]
                dist.all_gather(parts, mine)
                want = torch.stack(parts).mean(0)
                got = torch.view_as_real(blk.weight.grad)
                assert torch.allclose(got, want, rtol=1e-5, atol=1e-6), float((got - want).abs().max())
            assert not sync._pending
    finally:
        dist.destroy_process_group()


def test_chained_step_with_gradient_averaging_gloo():
    """bench.py at N > 1: the chained step's single backward pass fires the eight weight-gradient all-reduces from
    post-accumulate hooks (GradientAverager); world size 2, gloo, stand-in blocks."""
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_chained_dp_worker, args=(2, port), nprocs=2, join=True)


def _spatial_helpers_worker(rank, world, port):
    import sys
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        from benchlib import spatial
        h, polar, replica = spatial.make_groups(world, rank)
        assert h == world and replica == 0 and dist.get_world_size(group=polar) == world
        sizes = [3, 2]                                         # uneven shards, like 721 -> 181,181,181,178
        full = torch.arange(2 * 5 * 4, dtype=torch.float32).reshape(2, 5, 4)
        off = sum(sizes[:rank])
        got = spatial._gather_cat(full[:, off:off + sizes[rank]].contiguous(), 1, sizes, polar, h)
        assert torch.equal(got, full)
        cfull = torch.complex(full, -full)
        got = spatial._gather_cat(cfull[:, off:off + sizes[rank]].contiguous(), 1, sizes, polar, h)
        assert torch.equal(got, cfull)
    finally:
        dist.destroy_process_group()


def test_spatial_bench_helpers_gloo():
    """tools/benchlib/spatial.py: polar groups and the uneven gather used by the `parity` key of bench.py --gpus N."""
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_spatial_helpers_worker, args=(2, port), nprocs=2, join=True)
