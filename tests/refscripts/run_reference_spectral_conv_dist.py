"""One rank of: the reference's UNMODIFIED SpectralConv (makani/models/common/spectral_convolution.py:43-148) built on
the product's spatially sharded transforms.  The module reads `l_shapes / m_shapes / lat_shapes / lon_shapes` from the
inverse transform, checks `isinstance(..., thd.DistributedInverseRealSHT)` and asks `comm.get_rank("h" / "w")`
(:78-82) - all of that is answered by makani_b200.distributed and makani_b200.comm here.  The local stages of the
product transforms are CUDA-only, so they are replaced by the oracle's CPU arithmetic in test-only subclasses
(tests/test_distributed_cpu.py); groups, transposes, table slabs and shapes are the product code.
Compared with the single-process oracle block on the gathered field: output, input gradient, weight gradient (reduced
over "w", where dhconv weights are shared: is_shared_mp = ["matmul", "w"]).  Launched with torchrun (gloo)."""
import os
import sys
import warnings

warnings.filterwarnings("ignore")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import refharness as H                                                     # noqa: E402

H.install("product")
import torch                                                                # noqa: E402
import torch.distributed as dist                                            # noqa: E402
from makani_b200 import comm                                                # noqa: E402
from oracle import sht_oracle as O                                          # noqa: E402
from test_distributed_cpu import _make_cpu_classes                          # noqa: E402

sc = H.load("makani.models.common.spectral_convolution")
assert sc.__file__.startswith("/root/reference/")
h, w = int(os.environ["GRID_H"]), int(os.environ["GRID_W"])
comm.init(model_parallel_sizes=[h, w, 1, 1], model_parallel_names=["h", "w", "fin", "fout"])
thd, _CpuSHT, CpuISHT = _make_cpu_classes()


class CpuSHT(_CpuSHT):                      # keep the input precision (the reference module works in fp32 / complex64)
    def _dft_stage(self, x):
        return (2.0 * torch.pi * torch.fft.rfft(x, n=self.nlon, dim=-1, norm="forward"))[..., : self.mmax]


thd.init(comm.get_group("h") if h > 1 else None, comm.get_group("w") if w > 1 else None)
rh, rw = comm.get_rank("h"), comm.get_rank("w")

nlat, nlon, C, B, grid = 37, 50, 6, 2, "equiangular"
L, M = 18, 13
sht = CpuSHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
isht = CpuISHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
assert isinstance(isht, sc.thd.DistributedInverseRealSHT)
mod = sc.SpectralConv(sht, isht, C, C, operator_type="dhconv")
assert mod.modes_lat_local == sht.l_shapes[rh] and mod.modes_lon_local == sht.m_shapes[rw]
assert mod.nlat_local == sht.lat_shapes[rh] and mod.nlon_local == sht.lon_shapes[rw]
assert tuple(mod.weight.shape) == (C, C, sht.l_shapes[rh])

torch.manual_seed(3)                                                        # identical full tensors on every rank
x_full = torch.randn(B, C, nlat, nlon)
g_full = torch.randn(B, C, nlat, nlon)
w_full = torch.randn(C, C, L, dtype=torch.complex64) / C ** 0.5
sl = lambda t, dim, shapes, r: O.shard(t, dim, shapes, r)                    # noqa: E731
with torch.no_grad():
    mod.weight.copy_(sl(w_full, -1, sht.l_shapes, rh))
x = sl(sl(x_full, -2, sht.lat_shapes, rh), -1, sht.lon_shapes, rw).clone().requires_grad_(True)
y, res = mod(x)
assert res is x or torch.equal(res, x)                                      # same grid in and out: no re-projection
(y * sl(sl(g_full, -2, sht.lat_shapes, rh), -1, sht.lon_shapes, rw)).sum().backward()

ofwd = O.RealSHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
oinv = O.InverseRealSHT(nlat, nlon, lmax=L, mmax=M, grid=grid)
xr = x_full.double().requires_grad_(True)
wr = w_full.to(torch.complex128).requires_grad_(True)
yr, _ = O.spectral_conv_forward(xr, ofwd, oinv, wr, "dhconv", compute_dtype=torch.float64)
(yr * g_full.double()).sum().backward()

gw = mod.weight.grad.clone()
if w > 1:
    buf = torch.view_as_real(gw).contiguous()
    dist.all_reduce(buf, group=comm.get_group("w"))
    gw = torch.view_as_complex(buf)
errs = dict(y=O.rel_l2(y.detach(), sl(sl(yr.detach(), -2, sht.lat_shapes, rh), -1, sht.lon_shapes, rw)),
            dx=O.rel_l2(x.grad, sl(sl(xr.grad, -2, sht.lat_shapes, rh), -1, sht.lon_shapes, rw)),
            dw=float((gw - sl(wr.grad, -1, sht.l_shapes, rh)).norm() / sl(wr.grad, -1, sht.l_shapes, rh).norm()))
ok = all(e < 1e-5 for e in errs.values())                                   # the reference module computes in fp32
dist.barrier()
dist.destroy_process_group()
print(f"rank {comm.get_world_rank()}: reference SpectralConv on product distributed transforms {errs}: "
      f"{'OK' if ok else 'FAILED'}")
sys.exit(0 if ok else 1)
