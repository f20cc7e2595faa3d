"""Run in a fresh interpreter (tests/test_reference_dropin.py): the UNMODIFIED reference network
(makani/models/networks/sfnonet.py) is constructed on top of the PRODUCT modules - torch_harmonics shim,
makani_b200.SpectralConv in place of the reference's, makani_b200.comm as makani.utils.comm - exactly the
"zero-diff" route of INTEGRATION.md section 1.  Construction only: the product has no CPU path, and a forward on a
CPU tensor must say so loudly."""
import os
import sys
import warnings

warnings.filterwarnings("ignore")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import refharness as H                                                     # noqa: E402
import torch                                                               # noqa: E402

th = H.install("product")
import makani_b200                                                         # noqa: E402

sc = H.load("makani.models.common.spectral_convolution")
ref_spectral_conv = sc.SpectralConv
sc.SpectralConv = makani_b200.SpectralConv
common = H.load("makani.models.common")
common.SpectralConv = makani_b200.SpectralConv
net = H.load("makani.models.networks.sfnonet")
assert net.__file__.startswith("/root/reference/")
assert net.SpectralConv is makani_b200.SpectralConv and net.th.RealSHT is makani_b200.RealSHT

# config/sfnonet.yaml:24-40,174-187 at reduced size (same structure: sc3, 2 x resolution change, dhconv, instance norm)
kw = dict(inp_shape=(91, 180), out_shape=(91, 180), scale_factor=3, inp_chans=5, out_chans=4, embed_dim=16,
          num_layers=4, operator_type="dhconv", use_mlp=True, mlp_ratio=2, normalization_layer="instance_norm",
          big_skip=True, pos_embed="none", hard_thresholding_fraction=1.0, model_grid_type="equiangular",
          sht_grid_type="legendre-gauss")
torch.manual_seed(0)
model = net.SphericalFourierNeuralOperatorNet(**kw)

# the four transforms of sfnonet.py:536-539 are the product's, with the sizes _init_spectral_transforms derives
assert type(model.trans_down) is makani_b200.RealSHT and type(model.itrans_up) is makani_b200.InverseRealSHT
assert type(model.trans) is makani_b200.RealSHT and type(model.itrans) is makani_b200.InverseRealSHT
assert (model.trans_down.nlat, model.trans_down.nlon, model.trans_down.grid) == (91, 180, "equiangular")
assert (model.trans.nlat, model.trans.nlon, model.trans.grid) == (30, 60, "legendre-gauss")
L, M = 30, 31
for t in (model.trans_down, model.itrans_up, model.trans, model.itrans):
    assert (t.lmax, t.mmax) == (L, M), (t.lmax, t.mmax)
filters = [blk.filter.filter for blk in model.blocks]
assert len(filters) == 4 and all(type(f) is makani_b200.SpectralConv for f in filters)
assert filters[0].forward_transform is model.trans_down and filters[0].inverse_transform is model.itrans
assert filters[-1].forward_transform is model.trans and filters[-1].inverse_transform is model.itrans_up
assert [f.scale_residual for f in filters] == [True, False, False, True]
for f in filters:
    assert tuple(f.weight.shape) == (16, 16, L) and f.weight.dtype == torch.complex64
    assert f.weight.is_shared_mp == ["matmul", "w"] and f.weight.sharded_dims_mp == [None, None, "h"]

# checkpoint compatibility: identical parameter names / shapes / dtypes as the all-reference model (oracle-free: only
# the names matter), Legendre tables are not in the state dict on either side (trainer.py:1036-1042)
names = {k: (tuple(v.shape), v.dtype) for k, v in model.state_dict().items()}
assert not any("weights" in k.split(".")[-1] or k.endswith(".pct") for k in names), sorted(names)[:5]
sc.SpectralConv = ref_spectral_conv
common.SpectralConv = ref_spectral_conv
net.SpectralConv = ref_spectral_conv
torch.manual_seed(0)
model_ref = net.SphericalFourierNeuralOperatorNet(**kw)                     # reference SpectralConv on product SHT modules
names_ref = {k: (tuple(v.shape), v.dtype) for k, v in model_ref.state_dict().items()}
assert names == names_ref, set(names) ^ set(names_ref)
model.load_state_dict(model_ref.state_dict(), strict=True)

# the restated network that bench.py's `full_model` key times on the GPU box (tools/benchlib/sfno_model.py; the
# reference file cannot travel): same state-dict keys, shapes and dtypes as the reference network, strict load
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tools"))
from benchlib import sfno_model                                             # noqa: E402
restated = sfno_model.SFNO(inp_shape=(91, 180), scale_factor=3, inp_chans=5, out_chans=4, embed_dim=16, num_layers=4,
                           mlp_ratio=2)
names_restated = {k: (tuple(v.shape), v.dtype) for k, v in restated.state_dict().items()}
assert names_restated == names_ref, (set(names_restated) ^ set(names_ref))
restated.load_state_dict(model_ref.state_dict(), strict=True)
for a_, b_ in zip(restated.blocks, model_ref.blocks):                       # same initial scales (std within 25 %)
    for n_ in ("outer_skip.weight", "mlp.fwd.0.weight", "mlp.fwd.3.weight", "filter.filter.weight"):
        pa, pb = a_.get_parameter(n_), b_.get_parameter(n_)
        ra = float(torch.view_as_real(pa).std() if pa.is_complex() else pa.std())
        rb = float(torch.view_as_real(pb).std() if pb.is_complex() else pb.std())
        assert 0.75 < ra / rb < 1.33, (n_, ra, rb)

# block fusion on the reference's own FourierNeuralOperatorBlock instances: parameter names and
# objects unchanged, norm0 + GELU and norm1 + outer skip rewired, idempotent
before = {k: v.data_ptr() for k, v in model.state_dict().items()}
makani_b200.fuse_model(model)
makani_b200.fuse_model(model)
after = {k: v.data_ptr() for k, v in model.state_dict().items()}
assert before == after
for blk in model.blocks:
    assert type(blk).__name__ == "FourierNeuralOperatorBlock" and type(blk).__module__ == net.__name__
    assert isinstance(blk.norm0, makani_b200.FusedInstanceNorm2d) and blk.norm0.act == "gelu"
    assert isinstance(blk.act_layer0, torch.nn.Identity)
    assert isinstance(blk.norm1, makani_b200.FusedInstanceNorm2d) and blk.norm1.act == "none"
    assert blk._fused_outer_skip and not blk._fused_inner_skip and not hasattr(blk, "inner_skip")
    assert blk.norm0.eps == 1e-6 and blk.norm0.affine

# no CPU path: a forward on CPU tensors fails loudly instead of computing something else
try:
    model(torch.rand(1, 5, 91, 180))
except RuntimeError as e:
    assert "no CPU path" in str(e), e
else:
    raise AssertionError("forward on a CPU tensor did not raise")

# spectral_transform="fft" (FourierNeuralOperatorNet, sfnonet.py:541-553,648-650): the reference network on the
# product's RealFFT2 / InverseRealFFT2 (layers.py:217-280) and SpectralConv; same state-dict names as with its own
sc.SpectralConv = makani_b200.SpectralConv
common.SpectralConv = makani_b200.SpectralConv
net.SpectralConv = makani_b200.SpectralConv
ref_fft2 = (net.RealFFT2, net.InverseRealFFT2)
net.RealFFT2, net.InverseRealFFT2 = makani_b200.RealFFT2, makani_b200.InverseRealFFT2
torch.manual_seed(0)
fno = net.FourierNeuralOperatorNet(**kw)
assert type(fno.trans_down) is makani_b200.RealFFT2 and type(fno.itrans) is makani_b200.InverseRealFFT2
assert (fno.trans.lmax, fno.trans.mmax, fno.trans.truncate) == (30, 31, False)
assert (fno.trans_down.lmax, fno.trans_down.mmax, fno.trans_down.truncate) == (30, 31, True)
assert all(type(b.filter.filter) is makani_b200.SpectralConv for b in fno.blocks)
assert all(b.filter.filter._l_offset == makani_b200._ext.DENSE_L_OFFSET for b in fno.blocks)      # dense contraction
sc.SpectralConv = ref_spectral_conv
common.SpectralConv = ref_spectral_conv
net.SpectralConv = ref_spectral_conv
net.RealFFT2, net.InverseRealFFT2 = ref_fft2
torch.manual_seed(0)
fno_ref = net.FourierNeuralOperatorNet(**kw)
assert ({k: (tuple(v.shape), v.dtype) for k, v in fno.state_dict().items()}
        == {k: (tuple(v.shape), v.dtype) for k, v in fno_ref.state_dict().items()})
fno.load_state_dict(fno_ref.state_dict(), strict=True)
print("OK reference SFNO built on product modules:", len(names), "state-dict entries")
