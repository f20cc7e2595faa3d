"""Whole-model step for bench.py's `full_model` key (full-SFNO samples/s).

The reference network file (makani/models/networks/sfnonet.py) cannot travel to the GPU box and imports modulus /
torch-harmonics, which are absent; this module RESTATES its structure for the named configuration
(`sfno_linear_73chq_sc3_layers8_edim384_asgl2`: config/sfnonet.yaml:24-40,174-187) on the product modules, with the
reference's parameter names, shapes and initial scales, so that a reference checkpoint loads into it:

    encoder   EncoderDecoder, 1 hidden layer : Conv1x1(73 -> 384) + GELU + Conv1x1(384 -> 384, no bias)     layers.py:86-133
    8 blocks  filter (SpectralConv, dhconv) -> norm0 -> GELU -> MLP(384 -> 768 -> 384) -> norm1 -> + outer_skip(residual)
                                                                                                 sfnonet.py:119-267,386-431
    decoder   EncoderDecoder : Conv1x1(384 -> 384) + GELU + Conv1x1(384 -> 73, no bias)                      sfnonet.py:434-461
    big skip  out + Conv1x1(73 -> 73, no bias)(input)                                                        sfnonet.py:464-468,637-638

tests/refscripts/build_reference_sfno.py builds this class next to the UNMODIFIED reference network (in this container,
where /root/reference exists) and asserts that the two state dicts have identical keys, shapes and dtypes and that the
reference's state dict loads strictly.  As built, the 1x1 convolutions, norms and GELUs are torch modules around the
product's spectral path; makani_b200.fuse_model reroutes them (fused instance norms; 1x1-convolution stacks as library
GEMMs on the NCHW tensor with this repository's bias / GELU kernels in between).
"""
import math

import torch
import torch.nn as nn


def _conv1x1(cin, cout, bias, std):
    conv = nn.Conv2d(cin, cout, 1, bias=bias)
    nn.init.normal_(conv.weight, mean=0.0, std=std)
    if bias:
        nn.init.constant_(conv.bias, 0.0)
    return conv


class EncoderDecoder(nn.Module):
    """layers.py:86-133 (input_format="nchw"): `fwd` = [Conv1x1 + bias, act] * num_layers + Conv1x1 without bias."""

    def __init__(self, num_layers, input_dim, output_dim, hidden_dim, gain=1.0):
        super().__init__()
        mods, cur = [], input_dim
        for _ in range(num_layers):
            mods += [_conv1x1(cur, hidden_dim, True, math.sqrt(2.0 / cur)), nn.GELU()]
            cur = hidden_dim
        mods.append(_conv1x1(cur, output_dim, False, math.sqrt(gain / cur)))
        self.fwd = nn.Sequential(*mods)

    def forward(self, x):
        return self.fwd(x)


class MLP(nn.Module):
    """layers.py:136-214 without dropout: `fwd` = [fc1, GELU, Identity, fc2, Identity] (indices 0 and 3 carry weights)."""

    def __init__(self, in_features, hidden_features, gain=1.0):
        super().__init__()
        fc1 = _conv1x1(in_features, hidden_features, True, math.sqrt(2.0 / in_features))
        fc2 = _conv1x1(hidden_features, in_features, True, math.sqrt(gain / hidden_features))
        self.fwd = nn.Sequential(fc1, nn.GELU(), nn.Identity(), fc2, nn.Identity())

    def forward(self, x):
        return self.fwd(x)


class SpectralFilterLayer(nn.Module):
    """sfnonet.py:49-116, the branch filter_type="linear", factorization=None."""

    def __init__(self, forward_transform, inverse_transform, embed_dim, operator_type, gain):
        super().__init__()
        from makani_b200 import SpectralConv
        self.filter = SpectralConv(forward_transform, inverse_transform, embed_dim, embed_dim,
                                   operator_type=operator_type, separable=False, bias=False, gain=gain)

    def forward(self, x):
        return self.filter(x)


class Block(nn.Module):
    """sfnonet.py:119-267 with inner_skip="none", outer_skip="linear", use_mlp=True, no dropout, no final activation."""

    def __init__(self, forward_transform, inverse_transform, embed_dim, mlp_ratio, operator_type):
        super().__init__()
        norm = lambda: nn.InstanceNorm2d(embed_dim, eps=1e-6, affine=True, track_running_stats=False)   # noqa: E731
        self.norm0 = norm()
        self.filter = SpectralFilterLayer(forward_transform, inverse_transform, embed_dim, operator_type, gain=2.0)
        self.act_layer0 = nn.GELU()
        self.norm1 = norm()
        self.outer_skip = _conv1x1(embed_dim, embed_dim, False, math.sqrt(0.5 / embed_dim))
        self.mlp = MLP(embed_dim, int(embed_dim * mlp_ratio), gain=0.5)
        self.drop_path = nn.Identity()

    def forward(self, x):
        x, residual = self.filter(x)
        x = self.act_layer0(self.norm0(x))
        x = self.mlp(x)
        x = self.drop_path(self.norm1(x))
        return x + self.outer_skip(residual)


class SFNO(nn.Module):
    """sfnonet.py:270-640 for spectral_transform="sht", pos_embed="none", big_skip=True, inp_shape == out_shape."""

    def __init__(self, inp_shape=(721, 1440), scale_factor=3, inp_chans=73, out_chans=73, embed_dim=384, num_layers=8,
                 mlp_ratio=2, operator_type="dhconv", model_grid_type="equiangular", sht_grid_type="legendre-gauss",
                 hard_thresholding_fraction=1.0):
        super().__init__()
        import makani_b200 as mk
        self.inp_shape = self.out_shape = tuple(inp_shape)
        self.h, self.w = inp_shape[0] // scale_factor, inp_shape[1] // scale_factor
        lmax = int(self.h * hard_thresholding_fraction)
        mmax = int((self.w // 2 + 1) * hard_thresholding_fraction)
        self.trans_down = mk.RealSHT(*inp_shape, lmax=lmax, mmax=mmax, grid=model_grid_type).float()
        self.itrans_up = mk.InverseRealSHT(*inp_shape, lmax=lmax, mmax=mmax, grid=model_grid_type).float()
        self.trans = mk.RealSHT(self.h, self.w, lmax=lmax, mmax=mmax, grid=sht_grid_type).float()
        self.itrans = mk.InverseRealSHT(self.h, self.w, lmax=lmax, mmax=mmax, grid=sht_grid_type).float()
        self.encoder = EncoderDecoder(1, inp_chans, embed_dim, embed_dim)
        self.blocks = nn.ModuleList()
        for i in range(num_layers):
            fwd_t = self.trans_down if i == 0 else self.trans
            inv_t = self.itrans_up if i == num_layers - 1 else self.itrans
            self.blocks.append(Block(fwd_t, inv_t, embed_dim, mlp_ratio, operator_type))
        self.decoder = EncoderDecoder(1, embed_dim, out_chans, embed_dim, gain=0.5)
        self.residual_transform = _conv1x1(inp_chans, out_chans, False, math.sqrt(0.5 / inp_chans))

    def forward(self, x):
        residual = x
        x = self.encoder(x)
        for blk in self.blocks:
            x = blk(x)
        x = self.decoder(x)
        return x + self.residual_transform(residual)


class TorchSpectralConv(nn.Module):
    """The reference's SpectralConv.forward (spectral_convolution.py:124-148) on torch library calls: transforms with
    autocast disabled on x.float(), complex einsum, cast back - tools/benchlib/torch_gpu.py (cuFFT + cuBLAS)."""

    def __init__(self, weight, fwd, inv):
        super().__init__()
        self.weight = nn.Parameter(weight.detach().clone())
        self.fwd_t, self.inv_t = fwd, inv

    def forward(self, x):
        dtype = x.dtype
        residual = x
        x = x.float()
        with torch.autocast("cuda", enabled=False):
            x = self.fwd_t.sht(x).contiguous()
            if self.fwd_t is not self.inv_t:
                residual = self.inv_t.isht(x).to(dtype)
        x = torch.einsum("bixy,iox->boxy", x, self.weight).contiguous()
        with torch.autocast("cuda", enabled=False):
            x = self.inv_t.isht(x)
        return x.to(dtype), residual


def measure(dev, steps=5, warmup=2, batch=1):
    """One training step of the whole network at the named size: forward + backward under bf16 autocast
    (amp_mode bf16, trainer.py:516-537 `with amp.autocast(...)`), loss = mean squared output (a stand-in for the
    geometric L2 loss: one reduction over the output), gradients of all parameters; no optimizer update.  The same
    model instance is timed three times: as built (torch convolutions, norms and GELU around the product's spectral
    path), after fuse_model(pointwise=False) (fused instance norms), after fuse_model() (+ the 1x1-convolution stacks
    as GEMMs on the NCHW tensor with fused bias / GELU kernels), and with SpectralConv.enable_packed_master() on top."""
    import makani_b200 as mk
    from makani_b200 import _ext
    torch.manual_seed(0)
    model = SFNO().to(dev)
    nparams = sum(p.numel() * (2 if p.is_complex() else 1) for p in model.parameters())
    x = torch.randn(batch, 73, 721, 1440, device=dev)

    def step():
        for p in model.parameters():
            p.grad = None
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y = model(x)
            loss = (y.float() ** 2).mean()
        loss.backward()
        return loss

    def timed():
        for _ in range(warmup):
            step()
        torch.cuda.synchronize()
        torch.cuda.reset_peak_memory_stats()
        l0 = _ext.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            loss = step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        return {"ms_per_step": ms, "value": batch * 1e3 / ms, "unit": "samples/s", "steps": steps, "warmup": warmup,
                "own_launches_per_step": (_ext.launch_count() - l0) // steps,
                "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9, "loss": float(loss.detach()),
                "loss_finite": bool(torch.isfinite(loss).item())}

    out = {"batch": batch, "real_parameters": nparams}
    # the all-torch network first: what the unmodified reference runs on this GPU (cuDNN 1x1 convolutions, torch norms
    # and GELU, cuFFT + cuBLAS spectral path with TF32 allowed as makani/train.py:109-110 sets it)
    try:
        from . import torch_gpu
        tf32 = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
        torch.backends.cuda.matmul.allow_tf32 = torch.backends.cudnn.allow_tf32 = True
        L, M = model.trans.lmax, model.trans.mmax
        t_big = torch_gpu.TorchSHT(721, 1440, L, M, "equiangular", dev)
        t_small = torch_gpu.TorchSHT(model.h, model.w, L, M, "legendre-gauss", dev)
        own = [blk.filter.filter for blk in model.blocks]
        last = len(own) - 1
        for i, blk in enumerate(model.blocks):
            blk.filter.filter = TorchSpectralConv(own[i].weight, t_big if i == 0 else t_small, t_big if i == last else t_small)
        out["reference_equivalent_torch"] = timed()
        for blk, conv in zip(model.blocks, own):
            blk.filter.filter = conv
        del t_big, t_small
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = tf32
        torch.cuda.empty_cache()
    except Exception as e:                                      # noqa: BLE001
        out["reference_equivalent_torch"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    out["torch_glue"] = timed()
    mk.fuse_model(model, pointwise=False)
    out["fused_norm"] = timed()
    mk.fuse_model(model)
    out["fused"] = timed()
    for blk in model.blocks:                                   # opt-in: the spectral weights trained in the GEMMs' layout
        blk.filter.filter.enable_packed_master()
    out["fused_packed_master"] = timed()
    # the same step captured once and replayed as a CUDA graph (the reference's `fwdbwd` capture, trainer.py:84-152):
    # static input, gradients accumulate into static .grad buffers that are zeroed inside the graph
    try:
        for p in model.parameters():
            if p.requires_grad:
                p.grad = torch.zeros_like(p)

        def gstep():
            for p in model.parameters():
                if p.grad is not None:
                    p.grad.zero_()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                y = model(x)
                loss = (y.float() ** 2).mean()
            loss.backward()
            return loss

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                gstep()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            gloss = gstep()
        for _ in range(warmup):
            graph.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            graph.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out["fused_packed_master_graph"] = {"ms_per_step": ms, "value": batch * 1e3 / ms, "unit": "samples/s", "steps": steps,
                                            "loss": float(gloss.detach()), "loss_finite": bool(torch.isfinite(gloss).item())}
        del graph
    except Exception as e:                                      # noqa: BLE001
        out["fused_packed_master_graph"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    del model, x
    torch.cuda.empty_cache()
    return out
