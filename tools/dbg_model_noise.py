import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import makani_b200 as mk
from benchlib import sfno_model
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
DEV = "cuda:0"
torch.manual_seed(0)
net = sfno_model.SFNO(inp_shape=(45, 90), scale_factor=3, inp_chans=5, out_chans=5, embed_dim=16, num_layers=3).to(DEV)
x = torch.randn(2, 5, 45, 90, device=DEV)
def run(xx):
    for p in net.parameters(): p.grad = None
    y = net(xx); (y ** 2).mean().backward()
    return [y.detach().clone()] + [p.grad.clone() for p in net.parameters()]
def rel(a, b):
    a, b = (torch.view_as_real(a), torch.view_as_real(b)) if a.is_complex() else (a, b)
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))
names = ["y"] + [n for n, _ in net.named_parameters()]
w0 = run(x); w1 = run(x); wp = run(x * (1 + 1e-6))
print("repeat", max(rel(a, b) for a, b in zip(w1, w0)))
for n, a, b in zip(names, wp, w0):
    if rel(a, b) > 2e-4: print("perturb 1e-6", n, rel(a, b))
for stage, kw in (("norm", dict(pointwise=False)), ("all", {})):
    mk.fuse_model(net, **kw)
    g = run(x)
    for n, a, b in zip(names, g, w0):
        if rel(a, b) > 2e-4: print("fused", stage, n, rel(a, b))
