"""One rank of: the reference's UNMODIFIED DistributedInstanceNorm2d (makani/mpu/layer_norm.py:27-114) running on
the product's group registry (makani_b200.comm) and parallel-region mappings (makani_b200.distributed.mappings, in place
of modulus.distributed.mappings), compared with torch.nn.InstanceNorm2d on the gathered field - forward, input gradient
and affine-parameter gradients.  Spatial split h x w from GRID_H / GRID_W, uneven latitude / longitude slabs.
Launched by tests/test_reference_dropin.py with torchrun (gloo, CPU)."""
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
from makani_b200.distributed import compute_split_shapes                    # noqa: E402

ln = H.load("makani.mpu.layer_norm")
assert ln.__file__.startswith("/root/reference/")
h, w = int(os.environ["GRID_H"]), int(os.environ["GRID_W"])
comm.init(model_parallel_sizes=[h, w, 1, 1], model_parallel_names=["h", "w", "fin", "fout"])
rh, rw = comm.get_rank("h"), comm.get_rank("w")
assert comm.get_size("spatial") == h * w

torch.manual_seed(7)                                                        # same full field on every rank
B, C, K, W = 1, 5, 37, 50      # the reference reshapes its [B, C] statistics to [1, B*C, 1, 1] (layer_norm.py:89-90): B = 1 only
x_full = (torch.randn(B, C, K, W) * torch.linspace(0.5, 3.0, K).reshape(1, 1, K, 1) + 2.0).double()
g_full = torch.randn(B, C, K, W).double()
gamma, beta = torch.randn(C).double(), torch.randn(C).double()
lat, lon = compute_split_shapes(K, h), compute_split_shapes(W, w)


def local(t):
    return torch.split(torch.split(t, lat, dim=2)[rh], lon, dim=3)[rw].contiguous()


ref = torch.nn.InstanceNorm2d(C, eps=1e-5, affine=True).double()
with torch.no_grad():
    ref.weight.copy_(gamma)
    ref.bias.copy_(beta)
xr = x_full.clone().requires_grad_(True)
yr = ref(xr)
(yr * g_full).sum().backward()

mod = ln.DistributedInstanceNorm2d(C, eps=1e-5, affine=True).double()
assert mod.weight.is_shared_mp == ["spatial"]
with torch.no_grad():
    mod.weight.copy_(gamma)
    mod.bias.copy_(beta)
x = local(x_full).requires_grad_(True)
y = mod(x)
(y * local(g_full)).sum().backward()
# affine-parameter gradients are partial sums over this rank's slab: reduce over "spatial" (helpers / DDP do that)
gw, gb = mod.weight.grad.clone(), mod.bias.grad.clone()
dist.all_reduce(gw, group=comm.get_group("spatial"))
dist.all_reduce(gb, group=comm.get_group("spatial"))


def rel(a, b):
    return float((a - b).norm() / b.norm())


errs = dict(y=rel(y.detach(), local(yr.detach())), dx=rel(x.grad, local(xr.grad)), dgamma=rel(gw, ref.weight.grad),
            dbeta=rel(gb, ref.bias.grad))
ok = all(e < 1e-6 for e in errs.values())          # the reference computes its statistics in fp32 (layer_norm.py:94)
dist.barrier()
dist.destroy_process_group()
print(f"rank {comm.get_world_rank()}: reference DistributedInstanceNorm2d on product comm/mappings {errs}: "
      f"{'OK' if ok else 'FAILED'}")
sys.exit(0 if ok else 1)
