"""Per-stage device times of the h-sharded forward SHT at 721x1440 (peer-memory exchange).
   torchrun --nproc-per-node N tools/prof_spatial.py"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import makani_b200 as mk, makani_b200.distributed as thd
from makani_b200 import ops
from makani_b200.distributed import _PolarExchange

rank = int(os.environ["RANK"]); dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
torch.cuda.set_device(dev); dist.init_process_group("nccl", device_id=dev); thd.init(dist.group.WORLD, None)
C = 384
for (nlat, nlon, grid) in [(721, 1440, "equiangular"), (240, 480, "legendre-gauss")]:
    mod = thd.DistributedRealSHT(nlat, nlon, lmax=240, mmax=241, grid=grid).to(dev)
    x = torch.randn(1, C, mod.lat_shapes[rank], nlon, device=dev).to(torch.bfloat16)
    assert mod._peer_path(x)
    ex = _PolarExchange.of(mod, 1, C, dev); plan = mod._peer_plan(dev); sc = mod._scale(dev, "fwd")
    stages = {}
    fused = ex.dft_fused(mod)
    def run(record):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
        ev[0].record()
        if fused:
            ev[1].record(); xc = ex.dft_to_chan(x, sc)
        else:
            xm = ops.dft_fwd(x, mod.mmax, sc)
            ev[1].record(); xc = ex.lat_to_chan(xm)
        ev[2].record(); sp = ops.legendre_analysis(plan, xc, 1, ex.cs[ex.me])
        ev[3].record(); sl = ex.chan_to_l(sp)
        ev[4].record(); out = ops.sp_to_c64(sl, 1, C, ex.ls[ex.me], mod.mmax, ex.l0[ex.me], 0)
        ev[5].record()
        return ev
    for _ in range(3): run(False)
    torch.cuda.synchronize(); dist.barrier()
    acc = [0.0] * 5; n = 10
    t0 = time.perf_counter()
    evs = [run(True) for _ in range(n)]
    t_cpu = (time.perf_counter() - t0) / n * 1e3
    torch.cuda.synchronize()
    t_wall = (time.perf_counter() - t0) / n * 1e3
    for ev in evs:
        for i in range(5): acc[i] += ev[i].elapsed_time(ev[i + 1]) / n
    if rank == 0:
        print(json.dumps({"grid": f"{nlat}x{nlon}", "ranks": dist.get_world_size(), "fused_dft": fused, "dft_ms": acc[0], "lat_to_chan_ms (dft + exchange when fused)": acc[1],
                          "legendre_ms": acc[2], "chan_to_l_ms": acc[3], "sp_to_c64_ms": acc[4], "sum_ms": sum(acc),
                          "cpu_enqueue_ms_per_call": t_cpu, "wall_ms_per_call": t_wall}), flush=True)
dist.barrier(); dist.destroy_process_group()
