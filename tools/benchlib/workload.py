"""The benchmark workload: BASELINE.json configs[1], `sfno_linear_73chq_sc3_layers8_edim384_asgl2`
(config/sfnonet.yaml:174-187): the eight SpectralConv blocks of the model (sfnonet.py:392-393, 536-539)

    block 0      RealSHT(721x1440 equiangular) -> dhconv(384->384, L=240, M=241) -> InverseRealSHT(240x480 Gauss)
                 + residual re-projection to 240x480
    blocks 1..6  240x480 Gauss in and out
    block 7      240x480 in -> 721x1440 equiangular out (+ residual re-projection to 721x1440)

chained as inside the model, forward + backward, bf16 spatial tensors - plus the ALGORITHMIC byte counts the
rooflines are computed against."""
import torch

C_EMBED, L_MODES, M_MODES = 384, 240, 241
GRID_IN = (721, 1440, "equiangular")
GRID_INT = (240, 480, "legendre-gauss")
METRIC = "SFNO spectral path (8x SHT+SpectralConv+ISHT, fwd+bwd) samples/s"
WORKLOAD = "sfno_linear_73chq_sc3_layers8_edim384_asgl2: 8 SpectralConv blocks fwd+bwd, bf16 I/O"


def block_specs():
    """(in_grid, out_grid) of the eight blocks (sfnonet.py:392-393, 536-539)."""
    return [(GRID_IN, GRID_INT)] + [(GRID_INT, GRID_INT)] * 6 + [(GRID_INT, GRID_IN)]


# ------------------------------------------------------------------------------------------------ GPU arm
def build_gpu(dev, seed=0, channels=None):
    import makani_b200 as mk
    C = channels or C_EMBED
    torch.manual_seed(seed)
    sht_in = mk.RealSHT(*GRID_IN[:2], lmax=L_MODES, mmax=M_MODES, grid=GRID_IN[2]).to(dev)
    sht_int = mk.RealSHT(*GRID_INT[:2], lmax=L_MODES, mmax=M_MODES, grid=GRID_INT[2]).to(dev)
    isht_int = mk.InverseRealSHT(*GRID_INT[:2], lmax=L_MODES, mmax=M_MODES, grid=GRID_INT[2]).to(dev)
    isht_out = mk.InverseRealSHT(*GRID_IN[:2], lmax=L_MODES, mmax=M_MODES, grid=GRID_IN[2]).to(dev)
    blocks = []
    for gin, gout in block_specs():
        f = sht_in if gin is GRID_IN else sht_int
        i = isht_out if gout is GRID_IN else isht_int
        blocks.append(mk.SpectralConv(f, i, C, C, operator_type="dhconv").to(dev))
    return blocks


def make_inputs(batch, dtype, pinned, chained=False, channels=None):
    g = torch.Generator().manual_seed(1)
    xs = []
    for gin, _ in block_specs()[:1 if chained else None]:
        x = torch.rand(batch, channels or C_EMBED, gin[0], gin[1], generator=g, dtype=torch.float32).to(dtype)
        xs.append(x.pin_memory() if pinned else x)
    return xs


def make_grads(batch, dtype, dev, chained=False, channels=None):
    """Per block: (gradient of y, gradient of the re-projected residual or None).  Chained blocks get their output
    gradient from the next block, so only the last block's `gy` is materialised there."""
    g = torch.Generator().manual_seed(2)
    gs = []
    specs = block_specs()
    for i, (gin, gout) in enumerate(specs):
        rescale = (gin is not gout)
        if chained and i < len(specs) - 1 and not rescale:
            gs.append((None, None))
            continue
        gy = (torch.rand(batch, channels or C_EMBED, gout[0], gout[1], generator=g) - 0.5).to(dtype).to(dev)
        gs.append((gy, gy.clone() if rescale else None))
    return gs


def gpu_step(blocks, xs_dev, grads):
    """Every block forward + backward on its own input.  Returns a tiny checksum tensor."""
    sums = []
    for blk, x, (gy, gr) in zip(blocks, xs_dev, grads):
        x.grad = None
        blk.weight.grad = None
        y, res = blk(x)
        if gr is not None:
            torch.autograd.backward([y, res], [gy, gr])
        else:
            torch.autograd.backward([y], [gy])
        sums.append(y[0, 0, 0, :8].float().sum())
    return torch.stack(sums)


def gpu_step_chained(blocks, x0, grads):
    """One pass of the hot path as it runs inside the model (sfnonet.py:239-267 without the pointwise layers between
    the filters): the batch enters block 0, every block's output is the next block's input, and ONE backward pass
    carries the output gradient of block 7 (and the gradients of the two re-projected residuals) back to `x0`
    and to every weight.  The dhconv weight is re-packed by every forward (training mode), as after an optimizer
    update.  Returns a tiny checksum tensor."""
    x0.grad = None
    outs, gouts = [], []
    x = x0
    for blk, (gy, gr) in zip(blocks, grads):
        blk.weight.grad = None
        y, res = blk(x)
        if gr is not None:
            outs.append(res)
            gouts.append(gr)
        x = y
    outs.append(x)
    gouts.append(grads[-1][0])
    torch.autograd.backward(outs, gouts)
    return torch.stack([x[0, 0, 0, :8].float().sum(), x0.grad[0, 0, 0, :8].float().sum()])


def forward_chain(blocks, x):
    """Forward only (rollout, config 5): returns (output of block 7, its re-projected residual)."""
    res = None
    for blk in blocks:
        x, res = blk(x)
    return x, res


# ------------------------------------------------------------------------------------------------ algorithmic bytes
def tri(L, M):
    """number of (l, m <= l) pairs"""
    return sum(min(l + 1, M) for l in range(L))


def kernel_bytes(name, a):
    """Algorithmic HBM bytes of ONE launch of a kernel (fp32 internal tensors, dense table, the l >= m half of the
    spectral tensors) - formulas documented in DESIGN.md section 4."""
    if name in ("dft_fwd", "dft_inv"):
        return a["B"] * a["C"] * a["K"] * (a["W"] * a["esize"] + a["M"] * 8)
    if name in ("legendre_analysis", "legendre_synthesis"):
        # the table is charged to the launch that covers the first channel only (windows re-read it from L2)
        tab = 4 * a["M"] * a["L"] * a["K"] if a.get("c0", 0) == 0 else 0
        return a["B"] * a["C"] * 8 * (a["K"] * a["M"] + tri(a["L"], a["M"])) + tab
    if name in ("dhconv_fwd", "dhconv_bwd_x", "dhconv_bwd_w"):
        return 8 * a["B"] * (a["Cin"] + a["Cout"]) * tri(a["L"], a["M"]) + 8 * a["Cin"] * a["Cout"] * a["L"]
    if name in ("pack_weight", "unpack_wgrad", "unpack_wgrad_gather"):
        return 16 * a["Cin"] * a["Cout"] * a["L"]
    if name == "grad_slice_average":
        return 4 * a["n"] * (a["nsrc"] + 2)
    return 0


def field_bytes(grid, batch, C, esize):
    return batch * C * grid[0] * grid[1] * esize


def table_bytes(grid, L=L_MODES, M=M_MODES):
    return 4 * M * L * grid[0]


def block_forward_bytes(gin, gout, batch=1, C=C_EMBED, esize=2, L=L_MODES, M=M_MODES):
    """Algorithmic bytes of ONE fused SpectralConv block forward: field in + field out (+ the re-projected
    residual when the grids differ) + dhconv weight + the two Legendre tables.  esize = 4 gives SURVEY's fp32 figure
    (353.9 b + 394.2 MB for the internal block), esize = 2 the bf16-I/O variant (176.9 b + 394.2 MB)."""
    b = field_bytes(gin, batch, C, esize) + field_bytes(gout, batch, C, esize)
    if gin is not gout:
        b += field_bytes(gout, batch, C, esize)
    return b + 8 * C * C * L + table_bytes(gin, L, M) + table_bytes(gout, L, M)


def block_backward_bytes(gin, gout, batch=1, C=C_EMBED, esize=2, L=L_MODES, M=M_MODES):
    """Backward of one block: reads dy (+ d residual) and x, writes dx; reads W (for dX), writes dW; both tables."""
    b = field_bytes(gout, batch, C, esize) * (2 if gin is not gout else 1)
    b += 2 * field_bytes(gin, batch, C, esize)
    return b + 2 * 8 * C * C * L + table_bytes(gin, L, M) + table_bytes(gout, L, M)


def step_bytes(batch=1, C=C_EMBED, esize=2):
    """Algorithmic bytes of the whole step (8 blocks forward + backward); the weight re-pack / gradient unpack are
    layout changes, not algorithmic work, and count as zero."""
    return sum(block_forward_bytes(gi, go, batch, C, esize) + block_backward_bytes(gi, go, batch, C, esize)
               for gi, go in block_specs())
