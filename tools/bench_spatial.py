"""Timing of the h-sharded (polar) distributed transforms on one node - BASELINE.json config 4 (721x1440, h = 4).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 \
      tools/bench_spatial.py [--channels 384] [--steps 10]
  MAKANI_B200_P2P=0 ... (same)      # reference-style split / all_to_all / cat exchange for comparison

Per rank: x [1, C, 721/h, 1440] bf16 -> DistributedRealSHT -> [1, C, 240/h, 241] c64 -> DistributedInverseRealSHT
(240x480 Gauss grid) -> y; forward only and forward + backward.  Device time (CUDA events), max over ranks.
Prints one JSON line on rank 0."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import makani_b200 as mk
import makani_b200.distributed as thd


def timed(fn, steps, warmup, dev):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--channels", type=int, default=384)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    thd.init(dist.group.WORLD, None)
    C = args.channels
    sht_in = thd.DistributedRealSHT(721, 1440, lmax=240, mmax=241, grid="equiangular").to(dev)       # block 0 in
    isht_mid = thd.DistributedInverseRealSHT(240, 480, lmax=240, mmax=241, grid="legendre-gauss").to(dev)
    sht_mid = thd.DistributedRealSHT(240, 480, lmax=240, mmax=241, grid="legendre-gauss").to(dev)
    isht_out = thd.DistributedInverseRealSHT(721, 1440, lmax=240, mmax=241, grid="equiangular").to(dev)  # block 7 out
    torch.manual_seed(rank)
    x_hi = torch.randn(1, C, sht_in.lat_shapes[rank], 1440, device=dev).to(torch.bfloat16)
    x_mid = torch.randn(1, C, sht_mid.lat_shapes[rank], 480, device=dev).to(torch.bfloat16)
    out = {"config": {"workload": "distributed SHT/ISHT, h=%d w=1, B=1, C=%d, bf16 fields" % (world, C)},
           "n_gpus": world, "exchange": None}
    with torch.no_grad():
        c = sht_in(x_hi)
        out["sht_721x1440_ms"] = timed(lambda: sht_in(x_hi), args.steps, args.warmup, dev)
        out["isht_721x1440_ms"] = timed(lambda: isht_out(c), args.steps, args.warmup, dev)
        out["sht_240x480_ms"] = timed(lambda: sht_mid(x_mid), args.steps, args.warmup, dev)
        out["isht_240x480_ms"] = timed(lambda: isht_mid(c), args.steps, args.warmup, dev)

    xg = x_mid.float().requires_grad_(True)

    def fb():
        xg.grad = None
        y = isht_mid(sht_mid(xg))
        y.backward(y.detach())

    out["sht_isht_240x480_fwd_bwd_ms"] = timed(fb, args.steps, args.warmup, dev)
    st = thd.p2p.status()
    out["exchange"] = st["reason"] if st["enabled"] else "split / all_to_all / cat (%s)" % st["reason"]
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
