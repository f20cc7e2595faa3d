"""BASELINE.json config 4 (721x1440, h_parallel_size = 4; README recipe of the reference, consumer
makani/models/networks/sfnonet.py:528-533, makani/models/common/spectral_convolution.py:78-82) inside `bench.py --gpus N`:

* `parity`  - the h-sharded transforms and an h-sharded SpectralConv (block-0 geometry: 721x1440 -> 181,181,181,178
              latitudes per rank at h = 4) against the fp64 CPU oracle evaluated on rank 0, at C = 8: forward values,
              dX and dW, gathered over the ranks like the reference's tests/distributed/tests_fft.py:170-244 does.
              A mismatch makes the bench exit non-zero.
* `spatial` - ms per SpectralConv block forward + backward (block 0, internal, block 7) at C = 384, bf16, h ranks,
              next to the same block on one GPU, with the bytes every rank sends per block and the NVLink rate of the
              largest exchange alone.
N = 2 / 4: h = N.  N = 8: two data-parallel replicas of h = 4 (ranks 0-3 and 4-7)."""
import os

import torch
import torch.distributed as dist

from . import workload as W

TOL_T, TOL_BLOCK = 1e-3, 2e-3


def make_groups(world, rank):
    """Polar (h) groups of consecutive ranks; returns (h, my polar group, replica index)."""
    h = world if world <= 4 else 4
    mine = None
    for r0 in range(0, world, h):
        g = dist.new_group(list(range(r0, r0 + h)))
        if r0 <= rank < r0 + h:
            mine = g
    return h, mine, rank // h


def _gather_cat(t, dim, sizes, group, h):
    """Concatenate the ranks' shards (uneven along `dim`) on every rank of the group."""
    big = max(sizes)
    pad = list(t.shape)
    pad[dim] = big
    buf = torch.zeros(pad, dtype=t.dtype, device=t.device)
    buf.narrow(dim, 0, t.shape[dim]).copy_(t)
    if t.is_complex():
        outs = [torch.zeros_like(torch.view_as_real(buf)) for _ in range(h)]
        dist.all_gather(outs, torch.view_as_real(buf).contiguous(), group=group)
        outs = [torch.view_as_complex(o) for o in outs]
    else:
        outs = [torch.zeros_like(buf) for _ in range(h)]
        dist.all_gather(outs, buf.contiguous(), group=group)
    return torch.cat([o.narrow(dim, 0, n) for o, n in zip(outs, sizes)], dim=dim)


def _rel(a, b):
    a, b = a.double() if not a.is_complex() else a.to(torch.complex128), b
    return float((a.cpu() - b).norm() / b.norm().clamp_min(1e-300))


def parity(dev, rank, h, polar, replica, C=8):
    """Returns (dict of rel-L2 errors vs the fp64 oracle (world rank 0 only; None elsewhere), ok flag on every rank)."""
    import makani_b200 as mk
    import makani_b200.distributed as thd
    L, M = W.L_MODES, W.M_MODES
    gi, gt = W.GRID_IN, W.GRID_INT
    me = dist.get_rank(group=polar)
    dfwd = thd.DistributedRealSHT(gi[0], gi[1], lmax=L, mmax=M, grid=gi[2]).to(dev)
    dinv_big = thd.DistributedInverseRealSHT(gi[0], gi[1], lmax=L, mmax=M, grid=gi[2]).to(dev)
    dinv = thd.DistributedInverseRealSHT(gt[0], gt[1], lmax=L, mmax=M, grid=gt[2]).to(dev)
    assert dfwd.lat_shapes == thd.compute_split_shapes(gi[0], h)
    lat0, lat_int0 = sum(dfwd.lat_shapes[:me]), sum(dinv.lat_shapes[:me])
    l0 = sum(dfwd.l_shapes[:me])
    sl_lat = slice(lat0, lat0 + dfwd.lat_shapes[me])
    sl_lat_int = slice(lat_int0, lat_int0 + dinv.lat_shapes[me])
    sl_l = slice(l0, l0 + dfwd.l_shapes[me])

    g = torch.Generator().manual_seed(4321)
    x_full = torch.randn(1, C, gi[0], gi[1], generator=g)
    c_in = torch.view_as_complex(torch.randn(1, C, L, M, 2, generator=g)) * torch.tril(torch.ones(L, M))
    w_full = torch.view_as_complex(torch.randn(C, C, L, 2, generator=g)) / C ** 0.5
    gy_full = torch.randn(1, C, gt[0], gt[1], generator=g)
    gr_full = torch.randn(1, C, gt[0], gt[1], generator=g)

    # ---- distributed results, gathered
    x_loc = x_full[:, :, sl_lat].contiguous().to(dev)
    c_loc = dfwd(x_loc)
    c_all = _gather_cat(c_loc, 2, dfwd.l_shapes, polar, h)
    y_loc = dinv_big(c_in[:, :, sl_l].contiguous().to(torch.complex64).to(dev))
    y_all = _gather_cat(y_loc, 2, dinv_big.lat_shapes, polar, h)

    conv = mk.SpectralConv(dfwd, dinv, C, C, operator_type="dhconv").to(dev)
    with torch.no_grad():
        conv.weight.copy_(w_full[..., sl_l].to(torch.complex64))
    xg = x_loc.clone().requires_grad_(True)
    yc, rc = conv(xg)
    torch.autograd.backward([yc, rc], [gy_full[:, :, sl_lat_int].to(dev), gr_full[:, :, sl_lat_int].to(dev)])
    yc_all = _gather_cat(yc.detach(), 2, dinv.lat_shapes, polar, h)
    rc_all = _gather_cat(rc.detach(), 2, dinv.lat_shapes, polar, h)
    dx_all = _gather_cat(xg.grad, 2, dfwd.lat_shapes, polar, h)
    dw_all = _gather_cat(conv.weight.grad, 2, dfwd.l_shapes, polar, h)
    st = thd.p2p.status()
    peer = bool(st["enabled"]) and st["exchanges"] > 0
    torch.cuda.synchronize()

    errs = None
    ok = torch.ones(1, device=dev)
    if rank == 0:
        from oracle import sht_oracle as O
        osht = O.RealSHT(gi[0], gi[1], lmax=L, mmax=M, grid=gi[2])
        oisht_big = O.InverseRealSHT(gi[0], gi[1], lmax=L, mmax=M, grid=gi[2])
        oisht = O.InverseRealSHT(gt[0], gt[1], lmax=L, mmax=M, grid=gt[2])
        xr = x_full.double().requires_grad_(True)
        wr = w_full.to(torch.complex128).requires_grad_(True)
        errs = {f"h{h}_sht": _rel(c_all, osht(x_full.double())),
                f"h{h}_isht": _rel(y_all, oisht_big(c_in.to(torch.complex128)))}
        yr, rr = O.spectral_conv_forward(xr, osht, oisht, wr, "dhconv", compute_dtype=torch.float64)
        torch.autograd.backward([yr, rr], [gy_full.double(), gr_full.double()])
        errs.update({f"h{h}_conv_y": _rel(yc_all, yr.detach()), f"h{h}_conv_residual": _rel(rc_all, rr.detach()),
                     f"h{h}_conv_dx": _rel(dx_all, xr.grad), f"h{h}_conv_dw": _rel(dw_all, wr.grad)})
        bad = [k for k, v in errs.items() if not v < (TOL_BLOCK if "conv" in k else TOL_T)]
        errs["tolerance"] = {"transform": TOL_T, "block": TOL_BLOCK, "metric": "rel-L2 vs the fp64 CPU oracle"}
        errs["geometry"] = f"721x1440 -> lat {dfwd.lat_shapes}, l {dfwd.l_shapes}, C={C}, h={h}"
        errs["peer_memory_path"] = peer
        errs["exchange"] = st["reason"]
        if bad or not peer:
            errs["failed"] = bad + ([] if peer else ["peer_memory_path"])
            ok.zero_()
    dist.broadcast(ok, src=0)
    return errs, bool(ok.item() > 0)


def _time(fn, steps, warmup, dev, group):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def timings(dev, rank, h, polar, single_blocks, steps=8, warmup=2):
    """ms per block fwd+bwd on h ranks and on one GPU; `single_blocks` are the plain modules of the DP arm (the caller
    must have removed the gradient-averaging hooks from them: these runs are single-GPU work, no all-reduce)."""
    import makani_b200 as mk
    import makani_b200.distributed as thd
    C, L, M = W.C_EMBED, W.L_MODES, W.M_MODES
    gi, gt = W.GRID_IN, W.GRID_INT
    me = dist.get_rank(group=polar)
    dt = torch.bfloat16
    mods = {"sht_big": thd.DistributedRealSHT(gi[0], gi[1], lmax=L, mmax=M, grid=gi[2]).to(dev),
            "sht_int": thd.DistributedRealSHT(gt[0], gt[1], lmax=L, mmax=M, grid=gt[2]).to(dev),
            "isht_int": thd.DistributedInverseRealSHT(gt[0], gt[1], lmax=L, mmax=M, grid=gt[2]).to(dev),
            "isht_big": thd.DistributedInverseRealSHT(gi[0], gi[1], lmax=L, mmax=M, grid=gi[2]).to(dev)}
    kinds = {"block0": ("sht_big", "isht_int", 0), "internal": ("sht_int", "isht_int", 1), "block7": ("sht_int", "isht_big", 7)}
    out = {"h": h, "channels": C, "batch": 1, "io_dtype": "bf16", "blocks": {}}
    torch.manual_seed(100 + rank)
    for name, (f, i, idx) in kinds.items():
        fm, im = mods[f], mods[i]
        conv = mk.SpectralConv(fm, im, C, C, operator_type="dhconv").to(dev)
        x = torch.rand(1, C, fm.lat_shapes[me], fm.nlon, device=dev).to(dt).requires_grad_(True)
        gy = (torch.rand(1, C, im.lat_shapes[me], im.nlon, device=dev) - 0.5).to(dt)

        def step(conv=conv, x=x, gy=gy):
            x.grad = None
            conv.weight.grad = None
            y, res = conv(x)
            if conv.scale_residual:
                torch.autograd.backward([y, res], [gy, gy])
            else:
                torch.autograd.backward([y], [gy])

        ms_h = _time(step, steps, warmup, dev, polar)
        # the same block on ONE GPU (plain modules of the data-parallel arm; every rank runs it, no exchange)
        sb = single_blocks[idx]
        fs, is_ = sb.forward_transform, sb.inverse_transform
        xs = torch.rand(1, C, fs.nlat, fs.nlon, device=dev).to(dt).requires_grad_(True)
        gs = (torch.rand(1, C, is_.nlat, is_.nlon, device=dev) - 0.5).to(dt)

        def step1(sb=sb, xs=xs, gs=gs):
            xs.grad = None
            sb.weight.grad = None
            y, res = sb(xs)
            if sb.scale_residual:
                torch.autograd.backward([y, res], [gs, gs])
            else:
                torch.autograd.backward([y], [gs])

        ms_1 = _time(step1, steps, warmup, dev, polar)
        # bytes every rank sends per block fwd+bwd: each exchange moves (h-1)/h of the rank's local internal tensor
        fr = (h - 1) / h
        xm = lambda mod, Cc: M * 2 * Cc * (-(-mod.lat_shapes[me] // 4) * 4) * 4              # noqa: E731
        sp = lambda Cc: sum(min(M, (l + 16) // 16 * 16) for l in range(L)) * 2 * Cc * 4 / h      # noqa: E731  defined band, my l slab / channel slab
        n_syn = 2 if conv.scale_residual else 1
        sent = fr * ((xm(fm, C) + sp(C)) * 2                      # forward analysis + its adjoint (backward synthesis)
                     + (xm(im, C) + sp(C)) * 2 * n_syn)           # synthesis (+ residual) forward + adjoints
        out["blocks"][name] = {"ms_h": ms_h, "ms_1gpu": ms_1, "speedup": ms_1 / ms_h, "sent_MB_per_rank": sent / 1e6,
                               "nvlink_GBs_per_gpu_avg_over_block": sent / (ms_h * 1e-3) / 1e9}
        del conv, x, gy, xs, gs
    # the largest exchange alone: lat -> channel at 721x1440 (XM of the local latitude slab, all channels)
    from makani_b200 import ops
    fm = mods["sht_big"]
    ex = thd._PolarExchange.of(fm, 1, C, dev)
    x = torch.rand(1, C, fm.lat_shapes[me], fm.nlon, device=dev).to(dt)
    xm_loc = ops.dft_fwd(x, M, fm._scale(dev, "fwd"))
    ms_ex = _time(lambda: ex.lat_to_chan(xm_loc), steps, warmup, dev, polar)
    sent = (h - 1) / h * xm_loc.numel() * 4
    out["lat_to_chan_721"] = {"ms": ms_ex, "sent_MB_per_rank": sent / 1e6, "GBs_per_gpu": sent / (ms_ex * 1e-3) / 1e9,
                              "nvlink_peak_GBs": 900.0, "measured_peer_copy_GBs": 770.0}
    st = thd.p2p.status()
    out["exchange"] = st["reason"] if st["enabled"] else f"split / all_to_all / cat ({st['reason']})"
    return out
