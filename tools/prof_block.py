"""Small driver for ncu: a few SpectralConv passes at the SFNO shapes.
   python tools/prof_block.py [internal|block0|block7] [iters] [fwdbwd|fwd]
`fwd`: forward only in training mode (the weight re-pack launches every pass and can be told apart by kernel name)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import makani_b200 as mk

which = sys.argv[1] if len(sys.argv) > 1 else "internal"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
mode = sys.argv[3] if len(sys.argv) > 3 else "fwdbwd"
dev = torch.device("cuda:0")
C, L, M = 384, 240, 241
GI, GT = (721, 1440, "equiangular"), (240, 480, "legendre-gauss")
gin, gout = {"internal": (GT, GT), "block0": (GI, GT), "block7": (GT, GI)}[which]
fwd = mk.RealSHT(gin[0], gin[1], lmax=L, mmax=M, grid=gin[2]).to(dev)
inv = mk.InverseRealSHT(gout[0], gout[1], lmax=L, mmax=M, grid=gout[2]).to(dev)
conv = mk.SpectralConv(fwd, inv, C, C, operator_type="dhconv").to(dev)
xs = [torch.rand(1, C, gin[0], gin[1], device=dev).to(torch.bfloat16).requires_grad_(True) for _ in range(2)]
gy = (torch.rand(1, C, gout[0], gout[1], device=dev) - 0.5).to(torch.bfloat16)
for it in range(iters):
    x = xs[it % 2]
    x.grad = None
    conv.weight.grad = None
    y, res = conv(x)
    if mode == "fwdbwd":
        outs, gs = ([y, res], [gy, gy]) if conv.scale_residual else ([y], [gy])
        torch.autograd.backward(outs, gs)
    del y, res
torch.cuda.synchronize()
print("done", which, iters, mode, mk._ext.launch_count())
