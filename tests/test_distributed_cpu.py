"""World-size 2 and 4 gloo tests (CPU) of the spatially sharded SHT: split shapes, the all-to-all transposes
(forward and backward, uneven splits) and the full DistributedRealSHT / DistributedInverseRealSHT choreography.

The product's local stages are CUDA-only, so here they are replaced by the oracle's CPU arithmetic
(test-only subclass); everything else - process groups, transposes, table slabs, shapes - is the product code.
Pattern follows the reference's tests/distributed/tests_fft.py (local vs distributed, fwd and input-gradient,
rel-L2 <= 1e-6 after gather, uneven shapes such as 361 / 721-style splits).
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import sht_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _make_cpu_classes():
    import makani_b200.distributed as thd

    class CpuMixin:
        def _dft_stage(self, x):
            return (2.0 * torch.pi * torch.fft.rfft(x.double(), n=self.nlon, dim=-1, norm="forward"))[..., : self.mmax]

        def _idft_stage(self, c):
            return torch.fft.irfft(c, n=self.nlon, dim=-1, norm="forward")

        def _legendre_stage(self, c, synthesis):
            tab = (self.pct if synthesis else self.weights).to(c.real.dtype)     # local m-slab [Mloc, L, K]
            eq = "...lm,mlk->...km" if synthesis else "...km,mlk->...lm"
            return torch.complex(torch.einsum(eq, c.real, tab), torch.einsum(eq, c.imag, tab))

    class CpuSHT(CpuMixin, thd.DistributedRealSHT):
        pass

    class CpuISHT(CpuMixin, thd.DistributedInverseRealSHT):
        pass

    return thd, CpuSHT, CpuISHT


def _worker(rank, world, h, w, port, nlat, nlon, C, B):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        thd, CpuSHT, CpuISHT = _make_cpu_classes()
        # rank = rh * w + rw ; polar groups share rw, azimuth groups share rh
        rh, rw = rank // w, rank % w
        polar = azimuth = None
        for j in range(w):
            g = dist.new_group([i * w + j for i in range(h)])
            if j == rw:
                polar = g
        for i in range(h):
            g = dist.new_group([i * w + j for j in range(w)])
            if i == rh:
                azimuth = g
        thd.init(polar if h > 1 else None, azimuth if w > 1 else None)
        assert thd.polar_group_size() == h and thd.azimuth_group_size() == w
        assert thd.polar_group_rank() == rh and thd.azimuth_group_rank() == rw

        grid = "equiangular"
        L, M = nlat // 2, nlon // 4 + 1
        sht = CpuSHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
        isht = CpuISHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
        assert sht.lat_shapes == O.compute_split_shapes(nlat, h) and sht.lon_shapes == O.compute_split_shapes(nlon, w)
        assert sht.l_shapes == O.compute_split_shapes(L, h) and sht.m_shapes == O.compute_split_shapes(M, w)
        assert tuple(sht.weights.shape) == (sht.m_shapes[rw], L, nlat)

        # ---- transposes: forward == slice bookkeeping, backward == inverse transpose
        torch.manual_seed(0)
        full = torch.randn(B, C, nlat, nlon, dtype=torch.float64)
        loc = O.shard(O.shard(full, -2, sht.lat_shapes, rh), -1, sht.lon_shapes, rw).clone().requires_grad_(True)
        if w > 1:
            t = thd.distributed_transpose_azimuth.apply(loc, (1, -1), sht.lon_shapes)
            chans = O.compute_split_shapes(C, w)
            expect = O.shard(O.shard(full, -2, sht.lat_shapes, rh), 1, chans, rw)
            assert torch.equal(t, expect)
            back = thd.distributed_transpose_azimuth.apply(t, (-1, 1), chans)
            assert torch.equal(back, loc)
            back.square().sum().backward()
            assert torch.allclose(loc.grad, 2 * loc.detach())
            loc.grad = None

        # ---- full choreography vs the single-process oracle
        osht = O.RealSHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
        oisht = O.InverseRealSHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
        c_loc = sht(loc)
        c_exp = O.distributed_sht_expected(full, osht, h, w, rh, rw)
        assert c_loc.shape == c_exp.shape, (c_loc.shape, c_exp.shape)
        assert O.rel_l2(c_loc.detach(), c_exp) < 1e-6
        # input gradient of sum |c|^2 (global loss = sum over ranks): compare with the local autograd result
        fr = full.clone().requires_grad_(True)
        (osht(fr).abs() ** 2).sum().backward()
        (c_loc.abs() ** 2).sum().backward()
        g_exp = O.shard(O.shard(fr.grad, -2, sht.lat_shapes, rh), -1, sht.lon_shapes, rw)
        assert O.rel_l2(loc.grad, g_exp) < 1e-6

        cfull = osht(full)
        cl = O.shard(O.shard(cfull, -2, sht.l_shapes, rh), -1, sht.m_shapes, rw).clone().requires_grad_(True)
        x_loc = isht(cl)
        x_exp = O.distributed_isht_expected(cfull, oisht, h, w, rh, rw)
        assert x_loc.shape == x_exp.shape
        assert O.rel_l2(x_loc.detach(), x_exp) < 1e-6
        cr = cfull.clone().requires_grad_(True)
        oisht(cr).square().sum().backward()
        x_loc.square().sum().backward()
        gc_exp = O.shard(O.shard(cr.grad, -2, sht.l_shapes, rh), -1, sht.m_shapes, rw)
        assert O.rel_l2(cl.grad, gc_exp) < 1e-6

        # SpectralConv picks up the local shapes / offsets from the distributed inverse transform
        import makani_b200 as mk
        conv = mk.SpectralConv(sht, isht, C, C, operator_type="dhconv")
        assert tuple(conv.weight.shape) == (C, C, sht.l_shapes[rh])
        assert conv.modes_lon_local == sht.m_shapes[rw] and conv.nlat_local == sht.lat_shapes[rh]
        assert conv._l_offset == sum(sht.l_shapes[:rh]) and conv._m_offset == sum(sht.m_shapes[:rw])
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("h,w,nlat,nlon", [(2, 1, 37, 64), (1, 2, 24, 50), (2, 2, 37, 50)])
def test_distributed_sht_gloo(built_lib, h, w, nlat, nlon):
    world = h * w
    mp.spawn(_worker, args=(world, h, w, _free_port(), nlat, nlon, 6, 2), nprocs=world, join=True)


def test_split_helpers(built_lib):
    import makani_b200.distributed as thd
    assert thd.compute_split_shapes(721, 4) == [181, 181, 181, 178]
    assert thd.compute_split_shapes(5, 4) == [1, 1, 1, 2]
    for size in range(1, 40):
        for n in range(1, min(size, 8) + 1):
            assert thd.compute_split_shapes(size, n) == O.compute_split_shapes(size, n)
    parts = thd.split_tensor_along_dim(torch.arange(10), 0, 4)
    assert [p.numel() for p in parts] == [3, 3, 3, 1]
    with pytest.raises(ValueError):
        thd.split_tensor_along_dim(torch.arange(3), 0, 4)


# ------------------------------------------------------------------------------------------------ data parallel
def _dp_worker(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from makani_b200.distributed.grad_sync import GradientAverager
        torch.manual_seed(0)                                   # same parameters on every replica
        w = torch.nn.Parameter(torch.randn(3, 4, 5, dtype=torch.complex64))       # like SpectralConv.weight
        b = torch.nn.Parameter(torch.randn(4))
        w_ref = torch.nn.Parameter(w.detach().clone())         # same values, no hooks: the replica's own gradient
        b_ref = torch.nn.Parameter(b.detach().clone())
        sync = GradientAverager([w, b])
        gen = torch.Generator().manual_seed(100 + rank)        # different data per replica

        def loss(wp, bp, x):
            y = torch.einsum("bil,iol->bol", x.to(torch.complex64), wp) + bp[None, :, None]
            return (y.real ** 2 + y.imag ** 2).sum()

        for _ in range(2):                                      # two steps: hooks stay armed, nothing leaks between them
            x = torch.randn(2, 3, 5, generator=gen)
            for p in (w, b, w_ref, b_ref):
                p.grad = None
            loss(w, b, x).backward()                            # reductions start from the hooks during backward
            sync.wait()
            loss(w_ref, b_ref, x).backward()
            for got, mine in ((w.grad, w_ref.grad), (b.grad, b_ref.grad)):
                mine = torch.view_as_real(mine).contiguous() if mine.is_complex() else mine
                parts = [torch.zeros_like(mine) for _ in range(world)]
                dist.all_gather(parts, mine)
                want = torch.stack(parts).mean(0)
                got = torch.view_as_real(got) if got.is_complex() else got
                assert torch.allclose(got, want, rtol=1e-5, atol=1e-6), float((got - want).abs().max())
            assert not sync._pending and not sync._scaled
    finally:
        dist.destroy_process_group()


def test_data_parallel_gradient_averaging_gloo(built_lib):
    """The N > 1 path of bench.py: replicas + weight-gradient averaging started from post-accumulate hooks
    (makani_b200/distributed/grad_sync.py), complex parameters reduced through their real view."""
    mp.spawn(_dp_worker, args=(2, _free_port()), nprocs=2, join=True)
