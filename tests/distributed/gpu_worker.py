"""Multi-GPU parity worker (launched by torchrun from tests/test_gpu_distributed.py, one rank per GPU, NCCL).

Compares the h/w spatially sharded transforms (makani_b200.distributed) and a SpectralConv built on them with the
single-GPU modules on the same seeded inputs, forward and backward - the pattern of the reference's
tests/distributed/tests_fft.py:150-244 (local result -> split -> compare with the distributed result)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import torch.distributed as dist

import makani_b200 as mk
import makani_b200.distributed as thd

TOL = 2e-3     # relative L2; every transform is one TF32 contraction (tests/test_gpu_modules.py uses the same bound)


def rel(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def split_slice(shapes, idx):
    o = sum(shapes[:idx])
    return slice(o, o + shapes[idx])


def main():
    h, w = int(sys.argv[1]), int(sys.argv[2])
    nlat, nlon, grid, C = int(sys.argv[3]), int(sys.argv[4]), sys.argv[5], int(sys.argv[6])
    lmax = int(sys.argv[7]) if len(sys.argv) > 7 else None
    mmax = int(sys.argv[8]) if len(sys.argv) > 8 else None
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    assert h * w == world
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    ih, iw = rank // w, rank % w
    polar = azimuth = None
    for j in range(w):                       # polar (h) groups: same w index
        g = dist.new_group([i * w + j for i in range(h)])
        if j == iw:
            polar = g
    for i in range(h):                       # azimuth (w) groups: same h index
        g = dist.new_group([i * w + j for j in range(w)])
        if i == ih:
            azimuth = g
    thd.init(polar if h > 1 else None, azimuth if w > 1 else None)

    B = 2
    kw = dict(lmax=lmax, mmax=mmax, grid=grid)
    fwd = mk.RealSHT(nlat, nlon, **kw).to(dev)
    inv = mk.InverseRealSHT(nlat, nlon, **kw).to(dev)
    dfwd = thd.DistributedRealSHT(nlat, nlon, **kw).to(dev)
    dinv = thd.DistributedInverseRealSHT(nlat, nlon, **kw).to(dev)
    L, M = fwd.lmax, fwd.mmax
    assert dfwd.lat_shapes == thd.compute_split_shapes(nlat, h) and dfwd.m_shapes == thd.compute_split_shapes(M, w)
    sl_lat, sl_lon = split_slice(dfwd.lat_shapes, ih), split_slice(dfwd.lon_shapes, iw)
    sl_l, sl_m = split_slice(dfwd.l_shapes, ih), split_slice(dfwd.m_shapes, iw)

    g = torch.Generator(device="cpu").manual_seed(1234)
    x_full = torch.randn(B, C, nlat, nlon, generator=g).to(dev)
    gc_full = torch.randn(B, C, L, M, 2, generator=g).to(dev)
    gc_full = torch.view_as_complex(gc_full.contiguous())
    gx_full = torch.randn(B, C, nlat, nlon, generator=g).to(dev)
    errs = {}

    # ---- forward SHT, forward + backward
    xr = x_full.clone().requires_grad_(True)
    cr = fwd(xr)
    (cr * gc_full.conj()).real.sum().backward()
    xl = x_full[:, :, sl_lat, sl_lon].contiguous().requires_grad_(True)
    cl = dfwd(xl)
    (cl * gc_full[:, :, sl_l, sl_m].conj()).real.sum().backward()
    assert cl.shape == (B, C, dfwd.l_shapes[ih], dfwd.m_shapes[iw]), cl.shape
    errs["sht_fwd"] = rel(cl.detach(), cr.detach()[:, :, sl_l, sl_m])
    errs["sht_bwd"] = rel(xl.grad, xr.grad[:, :, sl_lat, sl_lon])

    # ---- inverse SHT, forward + backward (triangular coefficients, like real use)
    tri = (torch.arange(M, device=dev)[None, :] <= torch.arange(L, device=dev)[:, None])
    c_full = (gc_full * tri).contiguous()
    c1 = c_full.clone().requires_grad_(True)
    y1 = inv(c1)
    (y1 * gx_full).sum().backward()
    c2 = c_full[:, :, sl_l, sl_m].contiguous().requires_grad_(True)
    y2 = dinv(c2)
    (y2 * gx_full[:, :, sl_lat, sl_lon]).sum().backward()
    assert y2.shape == (B, C, dinv.lat_shapes[ih], dinv.lon_shapes[iw]), y2.shape
    errs["isht_fwd"] = rel(y2.detach(), y1.detach()[:, :, sl_lat, sl_lon])
    errs["isht_bwd"] = rel(c2.grad[..., :, :] * tri[sl_l, sl_m], (c1.grad * tri)[:, :, sl_l, sl_m])

    # ---- SpectralConv (dhconv) on the distributed transforms vs the fused single-GPU module
    torch.manual_seed(7)
    conv = mk.SpectralConv(fwd, inv, C, C, operator_type="dhconv").to(dev)
    dconv = mk.SpectralConv(dfwd, dinv, C, C, operator_type="dhconv").to(dev)
    wfull = conv.weight.detach().clone()
    dist.broadcast(torch.view_as_real(wfull), src=0)
    with torch.no_grad():
        conv.weight.copy_(wfull)
        dconv.weight.copy_(wfull[..., sl_l])
    assert dconv.weight.shape == (C, C, dfwd.l_shapes[ih])
    xr = x_full.clone().requires_grad_(True)
    yr, _ = conv(xr)
    (yr * gx_full).sum().backward()
    xl = x_full[:, :, sl_lat, sl_lon].contiguous().requires_grad_(True)
    yl, _ = dconv(xl)
    (yl * gx_full[:, :, sl_lat, sl_lon]).sum().backward()
    errs["conv_fwd"] = rel(yl.detach(), yr.detach()[:, :, sl_lat, sl_lon])
    errs["conv_dx"] = rel(xl.grad, xr.grad[:, :, sl_lat, sl_lon])
    # dhconv weights are shared over the azimuth group (is_shared_mp = ["matmul", "w"]): the reference reduces
    # their gradients over "w" (makani/mpu/mappings.py:104-172); do the same before comparing
    gw = dconv.weight.grad.clone()
    if w > 1:
        dist.all_reduce(torch.view_as_real(gw), group=azimuth)
    errs["conv_dw"] = rel(gw, conv.weight.grad[..., sl_l])

    # ---- spatially sharded fused instance norm (+ GELU) vs the single-GPU fused module on the full field
    # (statistics merged over all h x w ranks = the "spatial" group of the reference, mpu/layer_norm.py:47-86)
    torch.manual_seed(9)
    norm = mk.FusedInstanceNorm2d(C, eps=1e-6, affine=True, act="gelu").to(dev)
    with torch.no_grad():
        norm.weight.copy_(1 + 0.3 * torch.randn(C, device=dev))
        norm.bias.copy_(0.3 * torch.randn(C, device=dev))
        dist.broadcast(norm.weight, src=0)
        dist.broadcast(norm.bias, src=0)
    dnorm = mk.DistributedFusedInstanceNorm2d(C, eps=1e-6, affine=True, act="gelu", group=dist.group.WORLD).to(dev)
    dnorm.load_state_dict(norm.state_dict())
    xr = (x_full * 2 + 1).detach().requires_grad_(True)
    yr = norm(xr)
    (yr * gx_full).sum().backward()
    xl = (x_full * 2 + 1)[:, :, sl_lat, sl_lon].contiguous().requires_grad_(True)
    yl = dnorm(xl)
    (yl * gx_full[:, :, sl_lat, sl_lon]).sum().backward()
    gw, gb = dnorm.weight.grad.clone(), dnorm.bias.grad.clone()      # partial sums: shared over "spatial"
    dist.all_reduce(gw)
    dist.all_reduce(gb)
    norm_errs = dict(norm_fwd=rel(yl.detach(), yr.detach()[:, :, sl_lat, sl_lon]),
                     norm_dx=rel(xl.grad, xr.grad[:, :, sl_lat, sl_lon]),
                     norm_dw=rel(gw, norm.weight.grad), norm_db=rel(gb, norm.bias.grad))
    assert max(norm_errs.values()) < 1e-4, norm_errs                  # fp32 pointwise arithmetic, different summation order
    errs.update({k: v * (TOL / 1e-4) * 0.5 for k, v in norm_errs.items()})   # reported on the common scale below

    # ---- CUDA-graph replay of the h-sharded SpectralConv over peer memory (the reference captures forward + backward,
    # makani/utils/trainer.py:84-152): the barrier number lives on the device and every captured exchange is preceded
    # by a barrier, so replays with NEW inputs and ranks skewed in time must reproduce the eager results bit for bit.
    st = thd.p2p.status()
    if h > 1 and w == 1 and st["enabled"] and os.environ.get("MAKANI_TEST_SKIP_GRAPH", "0") != "1":
        xs = [torch.randn(B, C, dfwd.lat_shapes[ih], nlon, generator=torch.Generator().manual_seed(50 + k + 10 * rank)).to(dev)
              for k in range(3)]
        g_loc = gx_full[:, :, sl_lat, sl_lon].contiguous()
        x_static = xs[0].clone().requires_grad_(True)

        def gstep():
            x_static.grad = None
            dconv.weight.grad = None
            y, _ = dconv(x_static)
            torch.autograd.backward([y], [g_loc])
            return y

        # eager references on a side stream, as the reference's trainer warms up before capturing (trainer.py:109-130):
        # the leaves' gradient accumulators are bound to the stream of the forward pass that created them, and one that
        # is still alive and bound to the legacy default stream would drag that stream into the capture
        eager = []
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for k in range(3):
                with torch.no_grad():
                    x_static.copy_(xs[k])
                y = gstep()
                eager.append((y.detach().clone(), x_static.grad.clone(), dconv.weight.grad.clone()))
                del y
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        dist.barrier()
        graph = torch.cuda.CUDAGraph()
        x_static.grad = None
        dconv.weight.grad = None
        with torch.cuda.graph(graph):
            y_g = gstep()
        for k in (1, 2, 0, 2):
            with torch.no_grad():
                x_static.copy_(xs[k])
            torch.cuda._sleep(int(2e8) * (rank % 2 if k % 2 else (rank + 1) % 2))      # skew the ranks (~0.1 s)
            graph.replay()
            torch.cuda.synchronize()
            ye, gxe, gwe = eager[k]
            assert torch.equal(y_g, ye), ("graph replay: y differs", k)
            assert torch.equal(x_static.grad, gxe), ("graph replay: dx differs", k)
            assert torch.equal(torch.view_as_real(dconv.weight.grad), torch.view_as_real(gwe)), ("graph replay: dw", k)
        errs["graph_replay"] = 0.0

    # h-sharded runs must have gone through the peer-memory exchange unless it was switched off
    st = thd.p2p.status()
    if h > 1 and w == 1:
        want = os.environ.get("MAKANI_B200_P2P", "1") != "0"
        assert bool(st["enabled"]) == want and (st["exchanges"] > 0) == want, st

    torch.cuda.synchronize()
    worst = torch.tensor([max(errs.values())], device=dev)
    dist.all_reduce(worst, op=dist.ReduceOp.MAX)
    line = " ".join(f"{k}={v:.2e}" for k, v in errs.items())
    print(f"[rank {rank} h{ih} w{iw}] {line}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    if float(worst) >= TOL:
        print(f"DIST_GPU_FAIL worst={float(worst):.3e}", flush=True)
        sys.exit(1)
    if rank == 0:
        print(f"DIST_GPU_OK h={h} w={w} {nlat}x{nlon} worst={float(worst):.3e} exchange: {st['reason']} "
              f"({st['exchanges']} peer-memory scatters)", flush=True)


if __name__ == "__main__":
    main()
