"""h/w spatially sharded SHT / SpectralConv on real GPUs (NCCL all-to-all), one process per GPU via torchrun.
Skipped when fewer than two GPUs are visible (the CPU suite covers the same choreography with gloo)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "distributed", "gpu_worker.py")

CASES = [
    # h, w, nlat, nlon, grid, C, lmax, mmax, peer-memory exchange (h-sharded only; "0" forces the all-to-all path)
    (2, 1, 37, 72, "equiangular", 6, None, None, "1"),     # uneven latitude split 19/18, generic DFT kernel
    (2, 1, 37, 72, "equiangular", 5, None, None, "1"),     # uneven channel split 3/2 (padded channel slabs)
    (2, 1, 37, 72, "equiangular", 6, None, None, "0"),     # same through split / all_to_all / cat
    (2, 1, 60, 480, "legendre-gauss", 6, 60, 61, "1"),     # fft480 kernels: DFT fused with the exchange (tensor stores /
    (2, 1, 45, 1440, "equiangular", 5, 40, 41, "1"),       #   loads on peer memory), R = 1 and R = 3, uneven slabs 23/22, 3/2
    (1, 2, 32, 64, "legendre-gauss", 4, None, None, "1"),  # azimuth split: m 17/16
    (2, 2, 37, 72, "equiangular", 6, 30, 31, "1"),
    (4, 1, 240, 480, "legendre-gauss", 8, 240, 241, "1"),  # SFNO internal grid, fft480 kernels, l 60 x 4
    (4, 1, 721, 1440, "equiangular", 8, 240, 241, "1"),    # config 4: 721 -> 181,181,181,178
]


@pytest.mark.parametrize("h,w,nlat,nlon,grid,C,lmax,mmax,p2p", CASES)
def test_distributed_matches_single_gpu(built_lib, h, w, nlat, nlon, grid, C, lmax, mmax, p2p):
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n < h * w:
        pytest.skip(f"needs {h * w} GPUs, {n} visible")
    port = 29600 + (os.getpid() % 200)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={h * w}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), WORKER, str(h), str(w), str(nlat), str(nlon),
           grid, str(C)]
    if lmax is not None:
        cmd += [str(lmax), str(mmax)]
    env = dict(os.environ, MAKANI_B200_P2P=p2p, MAKANI_B200_P2P_FUSED_DFT="all")   # fused kernels at 480 and 1440
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    out = res.stdout + res.stderr
    if not (res.returncode == 0 and "DIST_GPU_OK" in out):
        keep = [ln for ln in out.splitlines() if any(k in ln for k in ("Error", "error", "rank", "assert", "DIST_GPU", "File "))]
        raise AssertionError("\n".join(keep[:60]) + "\n...\n" + out[-1500:])
