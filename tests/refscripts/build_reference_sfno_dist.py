"""One rank of: the UNMODIFIED reference network (sfnonet.py) constructed in SPATIALLY PARALLEL mode on the product
modules - `comm.get_size("spatial") > 1` makes `_init_spectral_transforms` (sfnonet.py:527-533,560-566) call
`thd.init(polar_group, azimuth_group)` with groups of makani_b200.comm and build
makani_b200.distributed.DistributedRealSHT / DistributedInverseRealSHT, read their `lat_shapes / lon_shapes`, and the
blocks pick the reference's DistributedInstanceNorm2d (on the product's mappings).  Construction only (no CPU path).
Launched with torchrun (gloo), GRID_H x GRID_W ranks."""
import os
import sys
import warnings

warnings.filterwarnings("ignore")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import refharness as H                                                     # noqa: E402

H.install("product")
import torch                                                                # noqa: E402
import torch.distributed as dist                                            # noqa: E402
import makani_b200                                                         # noqa: E402
import makani_b200.distributed as thd                                      # noqa: E402
from makani_b200 import comm                                                # noqa: E402

h, w = int(os.environ["GRID_H"]), int(os.environ["GRID_W"])
comm.init(model_parallel_sizes=[h, w, 1, 1], model_parallel_names=["h", "w", "fin", "fout"])
rh, rw = comm.get_rank("h"), comm.get_rank("w")

sc = H.load("makani.models.common.spectral_convolution")
sc.SpectralConv = makani_b200.SpectralConv
common = H.load("makani.models.common")
common.SpectralConv = makani_b200.SpectralConv
net = H.load("makani.models.networks.sfnonet")
assert net.__file__.startswith("/root/reference/") and net.SpectralConv is makani_b200.SpectralConv

# the 721 x 1440 / scale-factor-3 geometry of the named config (config/sfnonet.yaml:24-40,174-187) with few channels
# and a reduced mode count so that the four Legendre tables stay small on the CPU
torch.manual_seed(0)
model = net.SphericalFourierNeuralOperatorNet(
    inp_shape=(721, 1440), out_shape=(721, 1440), scale_factor=3, inp_chans=5, out_chans=4, embed_dim=8, num_layers=3,
    operator_type="dhconv", use_mlp=True, mlp_ratio=2, normalization_layer="instance_norm", big_skip=True,
    pos_embed="none", max_modes=(24, 25), model_grid_type="equiangular", sht_grid_type="legendre-gauss")

assert thd.is_initialized() and thd.polar_group_size() == h and thd.azimuth_group_size() == w
assert type(model.trans_down) is thd.DistributedRealSHT and type(model.itrans_up) is thd.DistributedInverseRealSHT
assert type(model.trans) is thd.DistributedRealSHT and type(model.itrans) is thd.DistributedInverseRealSHT
lat721, lon1440 = thd.compute_split_shapes(721, h), thd.compute_split_shapes(1440, w)
lat240, lon480 = thd.compute_split_shapes(240, h), thd.compute_split_shapes(480, w)
if h == 4:
    assert lat721 == [181, 181, 181, 178]
assert model.trans_down.lat_shapes == lat721 and model.trans_down.lon_shapes == lon1440
assert model.inp_shape_loc == (lat721[rh], lon1440[rw]) and model.out_shape_loc == (lat721[rh], lon1440[rw])
assert (model.h_loc, model.w_loc) == (lat240[rh], lon480[rw])
l_shapes, m_shapes = thd.compute_split_shapes(24, h), thd.compute_split_shapes(25, w)
for blk in model.blocks:
    f = blk.filter.filter
    assert type(f) is makani_b200.SpectralConv
    assert tuple(f.weight.shape) == (8, 8, l_shapes[rh]) and f.modes_lon_local == m_shapes[rw]
    assert type(blk.norm0).__name__ == "DistributedInstanceNorm2d" and type(blk.norm0).__module__ == "makani.mpu.layer_norm"
assert model.blocks[0].input_shape_loc == (lat721[rh], lon1440[rw])
assert model.blocks[0].output_shape_loc == (lat240[rh], lon480[rw])
# the Legendre tables every rank holds are its azimuth rank's m-slab (weights [M_loc, L, K])
assert tuple(model.trans.weights.shape) == (m_shapes[rw], 24, 240)

# block fusion recognises the reference's DistributedInstanceNorm2d
makani_b200.fuse_model(model)
for blk in model.blocks:
    assert isinstance(blk.norm0, makani_b200.DistributedFusedInstanceNorm2d) and blk.norm0.act == "gelu"
    assert blk.norm0._group() is comm.get_group("spatial")
    assert blk.norm0.weight.is_shared_mp == ["spatial"]
dist.barrier()
dist.destroy_process_group()
print(f"rank {comm.get_world_rank()}: reference SFNO (spatially parallel {h}x{w}) built on product modules: OK")
