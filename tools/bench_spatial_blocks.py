"""Config 4 alone: the `spatial` key of bench.py (h-sharded SpectralConv blocks fwd + bwd, the lat -> channel exchange)
without the data-parallel step around it - for A/B runs of the exchange, e.g.
    torchrun --nproc-per-node 4 tools/bench_spatial_blocks.py                      # SM row scatter (default)
    MAKANI_B200_P2P_COPY=ce torchrun --nproc-per-node 4 tools/bench_spatial_blocks.py   # copy-engine variant
Rank 0 prints one JSON line."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from benchlib import spatial, workload as W  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
h, polar, replica = spatial.make_groups(world, rank)
import makani_b200.distributed as thd  # noqa: E402
thd.init(polar, None)
blocks = W.build_gpu(dev)
out = spatial.timings(dev, rank, h, polar, blocks, steps=10, warmup=3)
out["copy"] = os.environ.get("MAKANI_B200_P2P_COPY", "sm")
if rank == 0:
    print(json.dumps(out), flush=True)
dist.barrier()
dist.destroy_process_group()
