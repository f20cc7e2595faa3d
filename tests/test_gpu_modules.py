"""GPU parity tests of the drop-in modules (RealSHT, InverseRealSHT, SpectralConv) against the CPU oracle
(oracle/sht_oracle.py, fp64) on the same seeded inputs, forward and backward, plus size-independent
properties at the full BASELINE.json sizes (240x480 / 721x1440, C = 384).

Metric: the reference's own (tests/distributed/tests_fft.py:231): mean over leading dims of
||a-b||_2 / ||b||_2 over the last two dims.  Stated tolerances vs the fp64 oracle:
   transforms (one TF32 contraction)            1e-3
   SpectralConv block (three TF32 contractions) 2e-3      gradients: same
   bf16 in/out                                  2e-2      (bf16 rounding of the output dominates)
"""
import pytest
import torch

from oracle import sht_oracle as O

pytestmark = pytest.mark.gpu

TOL_T = 1e-3
TOL_BLOCK = 2e-3


@pytest.fixture(scope="module")
def mk(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import makani_b200
    return makani_b200


DEV = "cuda:0"

GRIDS = [
    (32, 64, None, None, "legendre-gauss"),
    (33, 64, None, None, "equiangular"),        # odd nlat (latitude padding), poles included
    (17, 40, None, None, "lobatto"),
    (24, 48, 16, 12, "legendre-gauss"),         # truncated lmax / mmax
    (60, 480, 60, 61, "legendre-gauss"),        # fft480 path, R = 1
    (45, 1440, 40, 41, "equiangular"),          # fft480 path, R = 3
]


@pytest.mark.parametrize("nlat,nlon,lmax,mmax,grid", GRIDS)
def test_real_sht_forward_backward(mk, nlat, nlon, lmax, mmax, grid):
    torch.manual_seed(0)
    sht = mk.RealSHT(nlat, nlon, lmax=lmax, mmax=mmax, grid=grid).to(DEV)
    ref = O.RealSHT(nlat, nlon, lmax=lmax, mmax=mmax, grid=grid)
    x = torch.randn(2, 3, nlat, nlon, device=DEV, requires_grad=True)
    c = sht(x)
    assert c.shape == (2, 3, sht.lmax, sht.mmax) and c.dtype == torch.complex64
    xr = x.detach().double().cpu().requires_grad_(True)
    cr = ref(xr)
    assert O.rel_l2(c.detach().cpu(), cr.detach()) < TOL_T
    g = torch.randn_like(cr)
    (c * g.to(DEV).to(torch.complex64).conj()).real.sum().backward()
    (cr * g.conj()).real.sum().backward()
    assert O.rel_l2(x.grad.cpu(), xr.grad) < TOL_T


@pytest.mark.parametrize("nlat,nlon,lmax,mmax,grid", GRIDS)
def test_inverse_sht_forward_backward(mk, nlat, nlon, lmax, mmax, grid):
    torch.manual_seed(1)
    isht = mk.InverseRealSHT(nlat, nlon, lmax=lmax, mmax=mmax, grid=grid).to(DEV)
    ref = O.InverseRealSHT(nlat, nlon, lmax=lmax, mmax=mmax, grid=grid)
    L, M = isht.lmax, isht.mmax
    c0 = torch.randn(2, 3, L, M, dtype=torch.complex128) * torch.tril(torch.ones(L, M))
    c = c0.to(torch.complex64).to(DEV).requires_grad_(True)
    cr = c0.clone().requires_grad_(True)
    y, yr = isht(c), ref(cr)
    assert y.shape == (2, 3, nlat, nlon) and y.dtype == torch.float32
    assert O.rel_l2(y.detach().cpu(), yr.detach()) < TOL_T
    g = torch.randn_like(yr)
    (y * g.float().to(DEV)).sum().backward()
    (yr * g).sum().backward()
    mask = torch.tril(torch.ones(L, M))
    assert O.rel_l2(c.grad.cpu(), cr.grad * mask) < TOL_T


def _block(mk, nlat_in, nlon_in, grid_in, nlat_out, nlon_out, grid_out, L, M, Cin, Cout, op, B, dtype, bias=False):
    torch.manual_seed(2)
    fwd = mk.RealSHT(nlat_in, nlon_in, lmax=L, mmax=M, grid=grid_in).to(DEV)
    inv = mk.InverseRealSHT(nlat_out, nlon_out, lmax=L, mmax=M, grid=grid_out).to(DEV)
    conv = mk.SpectralConv(fwd, inv, Cin, Cout, operator_type=op, bias=bias).to(DEV)
    if bias:
        with torch.no_grad():
            conv.bias.normal_()
    ofwd = O.RealSHT(nlat_in, nlon_in, lmax=L, mmax=M, grid=grid_in)
    oinv = O.InverseRealSHT(nlat_out, nlon_out, lmax=L, mmax=M, grid=grid_out)
    x = torch.randn(B, Cin, nlat_in, nlon_in, device=DEV).to(dtype).requires_grad_(True)
    y, res = conv(x)
    xr = x.detach().double().cpu().requires_grad_(True)
    wr = conv.weight.detach().cpu().to(torch.complex128).requires_grad_(True)
    br = conv.bias.detach().double().cpu() if bias else None
    yr, resr = O.spectral_conv_forward(xr, ofwd, oinv, wr, op, bias=br, compute_dtype=torch.float64)
    tol = TOL_BLOCK if dtype == torch.float32 else 2e-2
    assert y.dtype == dtype and y.shape == yr.shape
    e_y = O.rel_l2(y.detach().float().cpu(), yr.detach())
    assert e_y < tol, ("y", e_y)
    if conv.scale_residual:
        e_r = O.rel_l2(res.detach().float().cpu(), resr.detach())
        assert e_r < tol, ("residual", e_r)
    else:
        assert res is x
    gy = torch.randn_like(yr)
    loss = (y.float() * gy.float().to(DEV)).sum()
    lossr = (yr * gy).sum()
    if conv.scale_residual:
        gr = torch.randn_like(resr)
        loss = loss + (res.float() * gr.float().to(DEV)).sum()
        lossr = lossr + (resr * gr).sum()
    loss.backward()
    lossr.backward()
    e_gx = O.rel_l2(x.grad.float().cpu(), xr.grad)
    e_gw = float((conv.weight.grad.cpu().to(torch.complex128) - wr.grad).norm() / wr.grad.norm())
    assert e_gx < tol, ("dx", e_gx)
    assert e_gw < tol, ("dw", e_gw)
    return e_y, e_gx, e_gw


@pytest.mark.parametrize("op", ["dhconv", "diagonal"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_spectral_conv_same_grid(mk, op, dtype):
    _block(mk, 32, 64, "legendre-gauss", 32, 64, "legendre-gauss", 32, 33, 12, 20, op, 2, dtype)


def test_spectral_conv_bias_and_residual_rescale(mk):
    # block-0 like: equiangular fine grid in, Gauss coarse grid out -> residual is re-projected
    _block(mk, 49, 96, "equiangular", 16, 32, "legendre-gauss", 16, 17, 8, 8, "dhconv", 2, torch.float32, bias="constant")
    # block-7 like: coarse in, fine out, position bias
    _block(mk, 16, 32, "legendre-gauss", 49, 96, "equiangular", 16, 17, 8, 8, "dhconv", 1, torch.float32, bias="position")


def test_spectral_conv_sfno_internal_shapes_small_channels(mk):
    # the internal SFNO grid (240x480 Gauss, L=240, M=241) with few channels so the fp64 oracle stays cheap
    _block(mk, 240, 480, "legendre-gauss", 240, 480, "legendre-gauss", 240, 241, 8, 8, "dhconv", 1, torch.float32)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_spectral_conv_sfno_block0_real_grids_small_channels(mk, dtype):
    """SFNO block 0 at its real geometry (sfnonet.py:392-393,536-539): 721x1440 equiangular in (K = 721 Legendre,
    `trans_down` table, nlon = 1440 DFT) -> L = 240, M = 241 -> 240x480 Gauss out, residual re-projected; few
    channels so the fp64 oracle stays cheap.  Output, residual, dX and dW against the oracle values."""
    _block(mk, 721, 1440, "equiangular", 240, 480, "legendre-gauss", 240, 241, 6, 8, "dhconv", 1, dtype)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_spectral_conv_sfno_block7_real_grids_small_channels(mk, dtype):
    """SFNO block 7: 240x480 Gauss in -> 721x1440 equiangular out (`itrans_up` table, K = 721 synthesis, inverse
    DFT at nlon = 1440), output and re-projected residual both at full resolution."""
    _block(mk, 240, 480, "legendre-gauss", 721, 1440, "equiangular", 240, 241, 8, 6, "dhconv", 1, dtype)


def test_state_dict_and_annotations(mk):
    fwd = mk.RealSHT(16, 32, grid="legendre-gauss")
    inv = mk.InverseRealSHT(16, 32, grid="legendre-gauss")
    conv = mk.SpectralConv(fwd, inv, 4, 6, operator_type="dhconv", bias="constant")
    sd = conv.state_dict()
    assert set(sd) == {"weight", "bias"}                       # tables are non-persistent buffers
    assert sd["weight"].dtype == torch.complex64 and tuple(sd["weight"].shape) == (4, 6, 16)
    assert conv.weight.is_shared_mp == ["matmul", "w"] and conv.weight.sharded_dims_mp == [None, None, "h"]
    assert mk.SpectralConvS2 is mk.SpectralConv
    d = mk.SpectralConv(fwd, inv, 4, 6, operator_type="diagonal")
    assert tuple(d.weight.shape) == (4, 6, 16, 17) and d.weight.sharded_dims_mp == [None, None, "h", "w"]
    with pytest.raises(ValueError):
        mk.SpectralConv(fwd, inv, 4, 6, operator_type="nope")


@pytest.mark.parametrize("case", ["internal", "down", "up"])
def test_channel_windows_are_bit_identical_to_the_single_pass(mk, case, monkeypatch):
    """The DFT <-> Legendre stages run window by window over the channels (ops.sht_to_sp / ops.sp_to_field) so that the
    Fourier coefficients stay in L2.  Same kernels, same arithmetic per output element: output, re-projected residual,
    dX, dW and the bias gradient must equal the single-pass path bit for bit (B = 2, three windows of 32 channels)."""
    from makani_b200 import ops
    gin, gout = {"internal": ((60, 480, "legendre-gauss"), (60, 480, "legendre-gauss")),
                 "down": ((45, 1440, "equiangular"), (60, 480, "legendre-gauss")),
                 "up": ((60, 480, "legendre-gauss"), (45, 1440, "equiangular"))}[case]
    L, M, C, B = 40, 41, 96, 2
    torch.manual_seed(11)
    fwd = mk.RealSHT(gin[0], gin[1], lmax=L, mmax=M, grid=gin[2]).to(DEV)
    inv = mk.InverseRealSHT(gout[0], gout[1], lmax=L, mmax=M, grid=gout[2]).to(DEV)
    conv = mk.SpectralConv(fwd, inv, C, C, operator_type="dhconv", bias="constant").to(DEV)
    with torch.no_grad():
        conv.bias.normal_()
    x = torch.randn(B, C, gin[0], gin[1], device=DEV).to(torch.bfloat16).requires_grad_(True)
    gy = torch.randn(B, C, gout[0], gout[1], device=DEV).to(torch.bfloat16)

    def run(window_mb):
        monkeypatch.setattr(ops, "_WINDOW_MB", window_mb)
        x.grad = conv.weight.grad = conv.bias.grad = None
        y, res = conv(x)
        outs, gs = ([y, res], [gy, gy]) if conv.scale_residual else ([y], [gy])
        torch.autograd.backward(outs, gs)
        return [t.detach().clone() for t in (y, res, x.grad, torch.view_as_real(conv.weight.grad), conv.bias.grad)]

    ref = run(0.0)
    assert ops.window_channels(C, B, gin[0], M) == C
    per_channel_mb = M * 2 * B * ((max(gin[0], gout[0]) + 3) // 4 * 4) * 4 / 1e6
    win = run(33 * per_channel_mb)                                   # room for 33 channels -> windows of 32
    assert ops.window_channels(C, B, max(gin[0], gout[0]), M) == 32
    for name, a, b in zip(["y", "residual", "dx", "dw", "dbias"], ref, win):
        assert torch.equal(a, b), name


def test_weight_update_behind_the_version_counter_is_seen(mk):
    """Optimizers that write through raw pointers or `.data` (apex FusedLAMB, selectable at makani/utils/trainer.py:458-470)
    do not bump `weight._version`; the packed copy of the dhconv weight must not go stale.  Training: re-packed every
    forward.  Inference: cached, `invalidate()` is the documented contract."""
    torch.manual_seed(5)
    fwd = mk.RealSHT(16, 32, grid="legendre-gauss").to(DEV)
    inv = mk.InverseRealSHT(16, 32, grid="legendre-gauss").to(DEV)
    conv = mk.SpectralConv(fwd, inv, 4, 4, operator_type="dhconv").to(DEV)
    x = torch.randn(1, 4, 16, 32, device=DEV)
    y0, _ = conv(x)
    v0 = conv.weight._version
    conv.weight.data.mul_(2.0)                               # in place behind autograd: version unchanged
    assert conv.weight._version == v0
    y1, _ = conv(x)
    assert O.rel_l2(y1.detach().cpu(), 2.0 * y0.detach().cpu()) < 1e-5
    with torch.no_grad():
        y2, _ = conv(x)                                      # fills the inference cache
        conv.weight.data.mul_(0.5)
        conv.invalidate()
        y3, _ = conv(x)
    assert O.rel_l2(y2.cpu(), y1.detach().cpu()) < 1e-6 and O.rel_l2(y3.cpu(), y0.detach().cpu()) < 1e-5


# ---------------------------------------------------------------- full BASELINE sizes: properties only
def test_full_size_roundtrip_and_linearity_internal_grid(mk):
    torch.manual_seed(3)
    nlat, nlon, L, M, C = 240, 480, 240, 241, 384
    sht = mk.RealSHT(nlat, nlon, lmax=L, mmax=M, grid="legendre-gauss").to(DEV)
    isht = mk.InverseRealSHT(nlat, nlon, lmax=L, mmax=M, grid="legendre-gauss").to(DEV)
    c = torch.randn(1, C, L, M, dtype=torch.complex64, device=DEV) * torch.tril(torch.ones(L, M, device=DEV))
    c[..., 0] = c[..., 0].real.to(torch.complex64)
    c[..., M - 1] = 0                                      # m = 240 has no degree l >= m when L = 240
    x = isht(c)
    c2 = sht(x)
    assert O.rel_l2(c2.cpu(), c.cpu()) < 2 * TOL_T          # sht(isht(c)) = c for band-limited c
    x2 = isht(c2)
    assert O.rel_l2(x2.cpu(), x.cpu()) < 2 * TOL_T          # isht o sht is a projection
    a, b = torch.randn_like(x), torch.randn_like(x)
    lin = sht(2.0 * a - 3.0 * b)
    assert O.rel_l2(lin.cpu(), (2.0 * sht(a) - 3.0 * sht(b)).cpu()) < 2 * TOL_T


def test_full_size_block0_shapes(mk):
    """721x1440 equiangular in -> 240x480 Gauss out at C = 384 (SFNO block 0): adjoint identity
    <conv(x), g> = <x, conv^T g> ties forward and backward together without an oracle at this size."""
    torch.manual_seed(4)
    C, L, M = 384, 240, 241
    fwd = mk.RealSHT(721, 1440, lmax=L, mmax=M, grid="equiangular").to(DEV)
    inv = mk.InverseRealSHT(240, 480, lmax=L, mmax=M, grid="legendre-gauss").to(DEV)
    conv = mk.SpectralConv(fwd, inv, C, C, operator_type="dhconv").to(DEV)
    x = torch.randn(1, C, 721, 1440, device=DEV, requires_grad=True)
    y, res = conv(x)
    assert y.shape == (1, C, 240, 480) and res.shape == (1, C, 240, 480)
    g, gr = torch.randn_like(y), torch.randn_like(res)
    ((y * g).sum() + (res * gr).sum()).backward()
    lhs = float((y.double() * g.double()).sum() + (res.double() * gr.double()).sum())
    rhs = float((x.detach().double() * x.grad.double()).sum())
    assert abs(lhs - rhs) / abs(lhs) < 5e-3, (lhs, rhs)
    assert torch.isfinite(conv.weight.grad.abs()).all()


def test_spectral_conv_cuda_graph_capture(mk):
    """The reference captures forward + backward of the model in a CUDA graph (makani/utils/trainer.py:84-152).
    The calls behind SpectralConv only enqueue work (no host sync, no allocation outside torch's caching allocator,
    tile-scheduler counters re-armed by the kernels themselves), so a warmed-up step must capture and replay
    bit-identically - on a side stream, as the trainer does."""
    torch.manual_seed(0)
    nlat, nlon, C = 60, 480, 24
    fwd = mk.RealSHT(nlat, nlon, lmax=60, mmax=61, grid="legendre-gauss").to(DEV)
    inv = mk.InverseRealSHT(nlat, nlon, lmax=60, mmax=61, grid="legendre-gauss").to(DEV)
    conv = mk.SpectralConv(fwd, inv, C, C, operator_type="dhconv").to(DEV)
    x = torch.randn(2, C, nlat, nlon, device=DEV).to(torch.bfloat16).requires_grad_(True)
    gy = torch.randn(2, C, nlat, nlon, device=DEV).to(torch.bfloat16)

    def step():
        x.grad = None
        conv.weight.grad = None
        y, _ = conv(x)
        torch.autograd.backward([y], [gy])
        return y

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):                       # warm-up: plans, tile lists, twiddle tables, allocator pools
            y_eager = step()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    gx_eager, gw_eager = x.grad.clone(), conv.weight.grad.clone()
    y_eager = y_eager.detach().clone()

    g = torch.cuda.CUDAGraph()
    x.grad = None
    conv.weight.grad = None
    with torch.cuda.graph(g):
        y_graph = step()
    for _ in range(3):                           # replays reuse the captured scheduler counters: they must re-arm
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(y_graph, y_eager)
    assert torch.equal(x.grad, gx_eager)
    assert torch.equal(torch.view_as_real(conv.weight.grad), torch.view_as_real(gw_eager))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_composite_c_abi_calls(mk, dtype):
    """sfno_sht_fwd / sfno_sht_inv / sfno_block_fwd (one C call per reference operator, caller-owned workspace)
    against the fp64 oracle and against the module path they must reproduce bit for bit."""
    from makani_b200 import ops
    torch.manual_seed(3)
    nlat, nlon, L, M, Cin, Cout, B = 60, 480, 60, 61, 12, 20, 2
    sht = mk.RealSHT(nlat, nlon, lmax=L, mmax=M, grid="legendre-gauss").to(DEV)
    isht = mk.InverseRealSHT(nlat, nlon, lmax=L, mmax=M, grid="legendre-gauss").to(DEV)
    conv = mk.SpectralConv(sht, isht, Cin, Cout, operator_type="dhconv", bias=False).to(DEV)
    x = torch.randn(B, Cin, nlat, nlon, device=DEV).to(dtype)
    pf, pi = sht._plan(x.device), isht._plan(x.device)
    tol_t = TOL_T
    tol_b = TOL_BLOCK if dtype == torch.float32 else 2e-2

    c = ops.sht_fwd(pf, x)
    c_mod = sht(x)
    assert torch.equal(torch.view_as_real(c), torch.view_as_real(c_mod))
    cr = O.RealSHT(nlat, nlon, lmax=L, mmax=M, grid="legendre-gauss")(x.double().cpu())
    assert O.rel_l2(c.cpu(), cr) < tol_t

    y = ops.sht_inv(pi, c_mod, torch.float32)
    assert torch.equal(y, isht(c_mod))
    yr = O.InverseRealSHT(nlat, nlon, lmax=L, mmax=M, grid="legendre-gauss")(cr)
    assert O.rel_l2(y.cpu(), yr) < tol_t

    wp = ops.pack_weight(conv.weight.detach())
    yb = ops.block_fwd(pf, pi, x, wp, Cout)
    yb_mod, _ = conv(x)
    assert yb.dtype == dtype and torch.equal(yb, yb_mod.detach())
    w64 = conv.weight.detach().cpu().to(torch.complex128)
    ybr = O.InverseRealSHT(nlat, nlon, lmax=L, mmax=M, grid="legendre-gauss")(torch.einsum("bixy,iox->boxy", cr, w64))
    assert O.rel_l2(yb.float().cpu(), ybr) < tol_b

    # a too small workspace is an error, not a silent overrun
    import ctypes as C
    from makani_b200 import _ext
    ws = torch.empty(1024, dtype=torch.uint8, device=DEV)
    out = torch.empty_like(c)
    rc = _ext.load().sfno_sht_fwd(pf.handle, C.c_void_p(x.data_ptr()), ops.dtype_code(x.dtype), B, Cin,
                                  C.c_void_p(out.data_ptr()), C.c_void_p(ws.data_ptr()), ws.numel(), None)
    assert rc != 0 and b"workspace" in _ext.load().sfno_last_error()


@pytest.mark.parametrize("name", ["internal", "down", "up", "diag"])
def test_spectral_conv_matches_reference_module_fixtures(mk, name):
    """makani_b200.SpectralConv against outputs and gradients recorded from the reference's own, unmodified
    SpectralConv (fp32, oracle SHT; tests/golden/spectral_conv_ref_*.npz, tools/make_golden.py).
    Tolerance: TOL_BLOCK (three TF32 contractions) on y, residual, dX, dW; bias gradient 1e-5 (a plain sum)."""
    from golden_cases import SPECTRAL_CONV_CASES, load
    gin, gout, L, M, B, Ci, Co, op, bias = SPECTRAL_CONV_CASES[name]
    z = load(name)
    fwd = mk.RealSHT(gin[0], gin[1], lmax=L, mmax=M, grid=gin[2]).float().to(DEV)
    inv = mk.InverseRealSHT(gout[0], gout[1], lmax=L, mmax=M, grid=gout[2]).float().to(DEV)
    conv = mk.SpectralConv(fwd, inv, Ci, Co, operator_type=op, bias=bias).to(DEV)
    assert tuple(conv.weight.shape) == z["weight"].shape and conv.scale_residual == bool(z["scale_residual"])
    with torch.no_grad():
        conv.weight.copy_(torch.from_numpy(z["weight"]))
        if bias:
            assert tuple(conv.bias.shape) == z["bias"].shape
            conv.bias.copy_(torch.from_numpy(z["bias"]))
    x = torch.from_numpy(z["x"]).to(DEV).requires_grad_(True)
    y, res = conv(x)
    assert y.dtype == torch.float32 and tuple(y.shape) == z["y"].shape
    e = O.rel_l2(y.detach().cpu(), torch.from_numpy(z["y"]))
    assert e < TOL_BLOCK, ("y", e)
    loss = (y * torch.from_numpy(z["gy"]).to(DEV)).sum()
    if conv.scale_residual:
        e = O.rel_l2(res.detach().cpu(), torch.from_numpy(z["residual"]))
        assert e < TOL_BLOCK, ("residual", e)
        loss = loss + (res * torch.from_numpy(z["gr"]).to(DEV)).sum()
    else:
        assert res is x
    loss.backward()
    e = O.rel_l2(x.grad.cpu(), torch.from_numpy(z["dx"]))
    assert e < TOL_BLOCK, ("dx", e)
    dw = torch.from_numpy(z["dweight"])
    e = float((conv.weight.grad.cpu() - dw).norm() / dw.norm())
    assert e < TOL_BLOCK, ("dweight", e)
    if bias:
        db = torch.from_numpy(z["dbias"])
        e = float((conv.bias.grad.cpu() - db).norm() / db.norm())
        assert e < 1e-5, ("dbias", e)


# ------------------------------------------------------------------------------------------------ packed master copy
def _small_conv(mk, seed=0):
    torch.manual_seed(seed)
    nlat, nlon, L, M, C = 24, 48, 24, 25, 12
    fwd = mk.RealSHT(nlat, nlon, lmax=L, mmax=M, grid="legendre-gauss").to(DEV)
    inv = mk.InverseRealSHT(nlat, nlon, lmax=L, mmax=M, grid="legendre-gauss").to(DEV)
    conv = mk.SpectralConv(fwd, inv, C, C, operator_type="dhconv", bias=False, gain=2.0).to(DEV)
    x = torch.randn(2, C, nlat, nlon, device=DEV)
    g = torch.randn(2, C, nlat, nlon, device=DEV)
    return conv, x, g


def test_packed_master_copy_matches_the_repacking_path(mk):
    """SpectralConv.enable_packed_master(): the trained parameter is the GEMM-layout tensor.  The weight gradient does
    not depend on the weight's rounding, so it must equal the default path's bit for bit (after the layout change);
    y and dx see the weight truncated instead of rounded to TF32: 1e-3 relative L2 (TOL_T)."""
    from makani_b200 import ops
    conv, x, g = _small_conv(mk)
    xa = x.clone().requires_grad_(True)
    ya, _ = conv(xa)
    ya.backward(g)
    gw_default = conv.weight.grad.clone()
    keys_default = {k: (tuple(v.shape), v.dtype) for k, v in conv.state_dict().items()}
    w0 = conv.weight.detach().clone()

    conv.enable_packed_master()
    assert conv.enable_packed_master() is conv                                  # idempotent
    assert not conv.weight.requires_grad and conv.weight_packed.requires_grad
    xb = x.clone().requires_grad_(True)
    yb, _ = conv(xb)
    yb.backward(g)
    Cin, Cout, L = conv.weight.shape
    gw_packed = ops.unpack_wgrad(conv.weight_packed.grad.reshape(-1), Cin, Cout, L)
    assert torch.equal(gw_packed, gw_default)
    assert O.rel_l2(yb.detach().cpu(), ya.detach().cpu()) < 1e-3
    assert O.rel_l2(xb.grad.cpu(), xa.grad.cpu()) < 1e-3

    # state dict: the reference's key / shape / dtype, exact values, no packed tensor in it
    sd = conv.state_dict()
    assert {k: (tuple(v.shape), v.dtype) for k, v in sd.items()} == keys_default
    assert torch.equal(sd["weight"], w0)


def test_packed_master_copy_optimizer_and_checkpoint_round_trip(mk):
    """One SGD step on the packed parameter equals the step on the complex parameter (elementwise on real numbers);
    a default module's state dict loads strictly into a packed module and gives the same forward."""
    conv_a, x, g = _small_conv(mk, seed=1)
    conv_b, _, _ = _small_conv(mk, seed=1)
    conv_b.load_state_dict(conv_a.state_dict(), strict=True)
    conv_b.enable_packed_master()
    for conv in (conv_a, conv_b):
        opt = torch.optim.SGD([p for p in conv.parameters() if p.requires_grad], lr=0.1, momentum=0.9)
        for _ in range(2):
            opt.zero_grad(set_to_none=True)
            y, _ = conv(x)
            y.backward(g)
            opt.step()
    wa, wb = conv_a.state_dict()["weight"], conv_b.state_dict()["weight"]
    assert O.rel_l2(torch.view_as_real(wb).cpu(), torch.view_as_real(wa).cpu()) < 2e-4   # the second step sees y through truncated weights

    conv_c, _, _ = _small_conv(mk, seed=2)
    conv_c.enable_packed_master()
    conv_c.load_state_dict(conv_a.state_dict(), strict=True)                    # default checkpoint -> packed module
    with torch.no_grad():
        ya, _ = conv_a(x)
        yc, _ = conv_c(x)
    assert O.rel_l2(yc.cpu(), ya.cpu()) < 1e-3
    assert torch.equal(conv_c.state_dict()["weight"], conv_a.state_dict()["weight"])
