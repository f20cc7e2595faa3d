"""GPU parity tests, kernel by kernel, through the C ABI (ops.* are 1:1 ctypes wrappers).
Each kernel is compared with an fp64 torch-CPU evaluation of the same contraction on seeded inputs.
Pad / undefined regions of the internal layouts are filled with NaN so any leak is visible.

Tolerances (relative L2 over the whole tensor):
  * DFT kernels (fp32 arithmetic):              1e-5
  * TF32 tensor-core contractions:              2e-3   (TF32 has a 10-bit mantissa; measured ~3e-4)
"""
import os

import pytest
import torch

from layouts import rel_err, rup, sp_decode, sp_encode, tf32_rn, tri_mask, wp_decode, xm_decode, xm_encode

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-5
TOL_TF32 = 2e-3


@pytest.fixture(scope="module")
def ops(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from makani_b200 import ops as _ops
    return _ops


def _dev():
    return torch.device("cuda:0")


def _rfft_ref(x64, M, scale):
    return (torch.fft.rfft(x64, dim=-1, norm="backward")[..., :M] * scale).to(torch.complex128)


@pytest.mark.parametrize("B,C,K,W,M,dtype", [
    (1, 3, 5, 16, 9, torch.float32),          # tiny, generic kernel, odd channel count (pad channels)
    (2, 4, 33, 64, 20, torch.float32),        # truncated bins, latitude padding (33 -> 36)
    (1, 5, 19, 90, 46, torch.bfloat16),
    (1, 8, 24, 480, 241, torch.float32),      # fft480 R=1
    (2, 3, 21, 480, 241, torch.bfloat16),
    (1, 4, 18, 1440, 241, torch.float32),     # fft480 R=3
    (1, 3, 17, 1440, 200, torch.float16),
    (1, 40, 240, 1440, 241, torch.bfloat16),  # 600 tiles: several tiles per CTA (dynamic scheduler, single row buffer)
    (2, 160, 240, 480, 241, torch.bfloat16),  # 4800 tiles at R=1 (double-buffered rows)
    (1, 3, 17, 1440, 721, torch.float32),     # more than 241 modes at nlon = 1440: three band launches (ops._dft_fwd_banded)
    (2, 5, 9, 1440, 600, torch.float32),
    (1, 4, 18, 1440, 400, torch.float32),
    (1, 2, 8, 1440, 481, torch.float32),
    (1, 2, 8, 1440, 480, torch.float32),
    (1, 2, 8, 1440, 242, torch.float32),
    (1, 4, 16, 1440, 721, torch.bfloat16),
])
def test_dft_fwd(ops, B, C, K, W, M, dtype):
    torch.manual_seed(0)
    x = torch.randn(B, C, K, W, device=_dev()).to(dtype)
    mscale = (torch.rand(W // 2 + 1, device=_dev()) + 0.5)
    xm = ops.dft_fwd(x, M, mscale, round_tf32=False)
    got = xm_decode(xm, B, C, K, M).cpu()
    ref = _rfft_ref(x.double().cpu(), M, mscale[:M].double().cpu())
    assert torch.isfinite(torch.view_as_real(got)).all()
    assert rel_err(got, ref) < TOL_FP32, rel_err(got, ref)
    # default: values pre-rounded to TF32 for the Legendre GEMM - exactly the RN rounding of the above
    got_r = xm_decode(ops.dft_fwd(x, M, mscale), B, C, K, M).cpu()
    assert torch.equal(got_r, tf32_rn(got))


@pytest.mark.parametrize("B,C,K,W,M,dtype", [
    (1, 3, 5, 16, 9, torch.float32),
    (2, 4, 33, 64, 20, torch.float32),
    (1, 5, 19, 90, 46, torch.float32),
    (1, 8, 24, 480, 241, torch.float32),
    (2, 3, 21, 480, 130, torch.float32),
    (1, 4, 18, 1440, 241, torch.float32),
    (1, 3, 17, 1440, 200, torch.float32),
    (1, 4, 16, 480, 241, torch.bfloat16),
    (1, 40, 240, 1440, 241, torch.bfloat16),  # 600 tiles: several tiles per CTA, single staging tile (NST = 1)
    (2, 160, 240, 480, 241, torch.bfloat16),  # 4800 tiles at R=1
    (1, 4, 18, 1440, 241, torch.float16),
    (1, 3, 17, 1440, 721, torch.float32),     # more than 241 modes at nlon = 1440: three band launches (ops._dft_inv_banded)
    (2, 5, 9, 1440, 600, torch.float32),
    (1, 4, 18, 1440, 400, torch.float32),
    (1, 2, 8, 1440, 481, torch.float32),
    (1, 2, 8, 1440, 480, torch.float32),
    (1, 2, 8, 1440, 242, torch.float32),
    (1, 4, 16, 1440, 721, torch.bfloat16),
])
def test_dft_inv(ops, B, C, K, W, M, dtype):
    torch.manual_seed(1)
    c = torch.randn(B, C, K, M, dtype=torch.complex64, device=_dev())
    mscale = (torch.rand(W // 2 + 1, device=_dev()) + 0.5)
    y = ops.dft_inv(xm_encode(c), B, C, K, W, M, mscale, dtype)
    cs = (c.cpu().to(torch.complex128) * mscale[:M].double().cpu())
    ref = torch.fft.irfft(cs, n=W, dim=-1, norm="forward")
    tol = TOL_FP32 if dtype == torch.float32 else 8e-3
    assert rel_err(y.cpu(), ref) < tol, rel_err(y.cpu(), ref)


def _table(M, L, K, seed):
    g = torch.Generator().manual_seed(seed)
    t = torch.randn(M, L, K, generator=g, dtype=torch.float64)
    return t * tri_mask(L, M, "cpu").t().unsqueeze(-1)      # zero for l < m, like the Legendre table


@pytest.mark.parametrize("B,C,K,L,M", [
    (1, 4, 32, 32, 33),
    (2, 3, 33, 20, 17),          # odd everything, pad channels, latitude pad
    (1, 70, 40, 40, 21),         # more than one 128-row tile (RS = 144)
    (1, 24, 300, 280, 100),      # > 256 degrees: two N tiles
    (1, 5, 721, 240, 241),       # the SFNO `trans_down` geometry: odd K = 721 (Kp = 724, TMA zero fill), L = 240, M = 241
    (1, 6, 240, 240, 241),       # the SFNO internal geometry
])
def test_legendre_analysis(ops, B, C, K, L, M):
    torch.manual_seed(2)
    tab = _table(M, L, K, 3)
    plan = ops.Plan(K, 2 * (M - 1), L, M, tab.numpy())
    x = torch.randn(B, C, K, M, dtype=torch.complex64, device=_dev())
    sp = ops.legendre_analysis(plan, xm_encode(x), B, C)
    got = sp_decode(sp, B, C, L, M).cpu()
    ref = torch.einsum("bckm,mlk->bclm", x.cpu().to(torch.complex128), tab.to(torch.complex128))
    assert torch.isfinite(torch.view_as_real(got)).all()
    assert rel_err(got, ref) < TOL_TF32, rel_err(got, ref)
    # the zero band m in (l, round_up(l+1,16)) must hold exact zeros (consumers rely on it)
    raw = sp.view(L, M, B, 2, rup(C, 4))[..., :C].cpu()
    for l in range(L):
        hi = min(M, (l + 16) // 16 * 16)
        assert (raw[l, l + 1:hi] == 0).all(), l


@pytest.mark.parametrize("B,C,K,L,M", [
    (1, 4, 32, 32, 33),
    (2, 3, 33, 20, 17),
    (1, 70, 140, 40, 21),        # two row tiles (K > 128), RS = 144
    (1, 136, 40, 24, 30),        # RS = 272 > 256: two column tiles; m >= L columns are all-zero
    (1, 5, 721, 240, 241),       # the SFNO `itrans_up` geometry: K = 721 -> six row tiles, the last one partial
    (1, 6, 240, 240, 241),       # the SFNO internal geometry
])
def test_legendre_synthesis(ops, B, C, K, L, M):
    torch.manual_seed(4)
    tab = _table(M, L, K, 5)
    plan = ops.Plan(K, 2 * (M - 1), L, M, tab.numpy())
    c = torch.randn(B, C, L, M, dtype=torch.complex64, device=_dev())
    xm = ops.legendre_synthesis(plan, sp_encode(c), B, C)
    got = xm_decode(xm, B, C, K, M).cpu()
    cm = c.cpu().to(torch.complex128) * tri_mask(L, M, "cpu")
    ref = torch.einsum("bclm,mlk->bckm", cm, tab.to(torch.complex128))
    assert torch.isfinite(torch.view_as_real(got)).all()
    assert rel_err(got, ref) < TOL_TF32, rel_err(got, ref)


@pytest.mark.parametrize("B,Cin,Cout,L,M", [
    (1, 8, 8, 20, 21),
    (2, 12, 20, 33, 17),
    (1, 130, 140, 40, 41),       # channel tiles > 128 on both sides
    (3, 16, 16, 150, 151),       # columns n = m*B + b exceed one 128-wide tile
    (1, 6, 10, 24, 13),          # channel counts that need padding (6 -> 8, 10 -> 12)
])
def test_dhconv(ops, B, Cin, Cout, L, M):
    torch.manual_seed(6)
    dev = _dev()
    x = torch.randn(B, Cin, L, M, dtype=torch.complex64, device=dev)
    w = torch.randn(Cin, Cout, L, dtype=torch.complex64, device=dev) / Cin ** 0.5
    g = torch.randn(B, Cout, L, M, dtype=torch.complex64, device=dev)
    tri = tri_mask(L, M, "cpu")
    x64 = x.cpu().to(torch.complex128) * tri
    g64 = g.cpu().to(torch.complex128) * tri
    w64 = w.cpu().to(torch.complex128)

    wp = ops.pack_weight(w)
    assert torch.equal(wp_decode(wp, Cin, Cout, L), tf32_rn(w))            # packed copy is TF32-rounded
    assert torch.equal(ops.unpack_wgrad(wp, Cin, Cout, L), tf32_rn(w))

    y = sp_decode(ops.dhconv_fwd(sp_encode(x), wp, B, Cin, Cout, L, M), B, Cout, L, M).cpu()
    ref = torch.einsum("bixy,iox->boxy", x64, w64)
    assert rel_err(y, ref) < TOL_TF32, ("fwd", rel_err(y, ref))

    gx = sp_decode(ops.dhconv_bwd_x(sp_encode(g), wp, B, Cin, Cout, L, M), B, Cin, L, M).cpu()
    ref = torch.einsum("boxy,iox->bixy", g64, w64.conj()) * tri
    assert rel_err(gx, ref) < TOL_TF32, ("bwd_x", rel_err(gx, ref))

    gw = ops.unpack_wgrad(ops.dhconv_bwd_w(sp_encode(x), sp_encode(g), B, Cin, Cout, L, M), Cin, Cout, L).cpu()
    ref = torch.einsum("bixy,boxy->iox", x64.conj(), g64)
    assert rel_err(gw, ref) < TOL_TF32, ("bwd_w", rel_err(gw, ref))


@pytest.mark.parametrize("B,Cin,Cout,L,M", [
    (2, 6, 10, 20, 21),
    (1, 16, 16, 40, 33),
    (2, 256, 256, 8, 9),         # embed_dim 256 at local batch 2 (staged tiles must be chunked over channels)
    (5, 40, 36, 12, 13),         # batch walked in chunks of 4 + 1
    (1, 420, 404, 6, 7),         # C >= 400 at batch 1
])
def test_diagonal(ops, B, Cin, Cout, L, M):
    torch.manual_seed(7)
    dev = _dev()
    x = torch.randn(B, Cin, L, M, dtype=torch.complex64, device=dev)
    w = torch.randn(Cin, Cout, L, M, dtype=torch.complex64, device=dev)
    g = torch.randn(B, Cout, L, M, dtype=torch.complex64, device=dev)
    tri = tri_mask(L, M, "cpu")
    x64, g64, w64 = x.cpu().to(torch.complex128) * tri, g.cpu().to(torch.complex128) * tri, w.cpu().to(torch.complex128)
    y = sp_decode(ops.diag_fwd(sp_encode(x), w, B, Cin, Cout, L, M), B, Cout, L, M).cpu()
    assert rel_err(y, torch.einsum("bixy,ioxy->boxy", x64, w64)) < TOL_FP32
    gx = sp_decode(ops.diag_bwd_x(sp_encode(g), w, B, Cin, Cout, L, M), B, Cin, L, M).cpu()
    assert rel_err(gx, torch.einsum("boxy,ioxy->bixy", g64, w64.conj()) * tri) < TOL_FP32
    gw = ops.diag_bwd_w(sp_encode(x), sp_encode(g), B, Cin, Cout, L, M).cpu()
    assert rel_err(gw, torch.einsum("bixy,boxy->ioxy", x64.conj(), g64)) < TOL_FP32


@pytest.mark.parametrize("B,C,L,M", [(1, 3, 20, 21), (2, 37, 33, 40)])
def test_layout_roundtrips(ops, B, C, L, M):
    torch.manual_seed(8)
    c = torch.randn(B, C, L, M, dtype=torch.complex64, device=_dev())
    cr = tf32_rn(c)                                        # conversions INTO the GEMM layouts round to TF32
    sp = ops.c64_to_sp(c)
    assert torch.equal(sp_decode(sp, B, C, L, M), cr * tri_mask(L, M, c.device))
    assert torch.equal(ops.sp_to_c64(sp, B, C, L, M), cr * tri_mask(L, M, c.device))
    assert torch.equal(ops.sp_to_c64(sp_encode(c), B, C, L, M), c * tri_mask(L, M, c.device))   # out of them: exact
    xm = ops.c64_to_xm(c)
    assert torch.equal(xm_decode(xm, B, C, L, M), cr)
    assert torch.equal(ops.xm_to_c64(xm, B, C, L, M), cr)
    assert torch.equal(ops.xm_to_c64(xm_encode(c), B, C, L, M), c)


def test_bias(ops):
    y = torch.randn(2, 5, 7, 9, device=_dev())
    b0 = torch.randn(1, 5, 1, 1, device=_dev())
    b1 = torch.randn(1, 5, 7, 9, device=_dev())
    assert torch.allclose(ops.add_bias_(y.clone(), b0), y + b0)
    assert torch.allclose(ops.add_bias_(y.clone(), b1), y + b1)


# ------------------------------------------------------------------------------------------------ peer-memory exchange
@pytest.mark.parametrize("h,nlat,L,M,B,C", [(2, 37, 20, 21, 1, 6), (3, 23, 17, 9, 2, 7), (4, 48, 33, 34, 1, 16)])
def test_rows_scatter_and_peer_barrier_single_gpu(built_lib, monkeypatch, h, nlat, L, M, B, C):
    """The exchange kernels of the h-sharded transforms (sfno_rows_scatter into sfno_ipc_alloc'ed buffers, then
    sfno_peer_barrier) on ONE GPU: the h ranks are played one after the other by this process, every 'peer' buffer is
    a buffer of this device.  Same checks as the CPU emulation (tests/test_exchange_cpu.py); the multi-GPU runs over
    NVLink are in tests/test_gpu_distributed.py.  (4, 48, ..., 16): 16-byte aligned everywhere -> the vector kernel.)"""
    import ctypes as C_
    import makani_b200.distributed as thd
    from makani_b200 import _ext
    from makani_b200.distributed import p2p
    from test_exchange_cpu import check_exchange
    lib = _ext.load()
    dev = torch.device("cuda:0")
    state = {"rank": 0, "h": h}
    pool = {}

    class DevBuffers:
        def __init__(self, key, nfloats):
            self._use = 0
            if (key, nfloats) not in pool:
                bufs = []
                for b in range(2):
                    ptrs = []
                    for r in range(h):
                        p, hnd = C_.c_void_p(), C_.create_string_buffer(64)
                        _ext.check(lib.sfno_ipc_alloc(nfloats * 4, C_.byref(p), hnd), "sfno_ipc_alloc")
                        ptrs.append(p.value)
                    bufs.append(ptrs)
                pool[(key, nfloats)] = bufs
            self.peer_ptrs = pool[(key, nfloats)]
            self.local = [torch.as_tensor(p2p._DevPtr(self.peer_ptrs[b][state["rank"]], nfloats), device=dev)
                          for b in range(2)]

        def next(self):
            b = self._use & 1
            self._use += 1
            return b

    flags, hnd = C_.c_void_p(), C_.create_string_buffer(64)
    _ext.check(lib.sfno_ipc_alloc(256, C_.byref(flags), hnd), "sfno_ipc_alloc")
    epoch = [0]

    class Barrier:                       # size-1 barrier with the barrier number kept on the device (epoch argument 0)
        def __call__(self):
            epoch[0] += 1
            arr = (C_.c_void_p * 1)(flags.value)
            stream = C_.c_void_p(torch.cuda.current_stream().cuda_stream)
            _ext.check(lib.sfno_peer_barrier(C_.cast(arr, C_.c_void_p), 0, 1, 0, stream), "sfno_peer_barrier")

        def pre_write(self):
            pass

    barrier = Barrier()
    monkeypatch.setattr(p2p, "buffers", lambda key, n, group, device: DevBuffers(key, n))
    monkeypatch.setattr(p2p, "fence", lambda group, device: barrier)
    monkeypatch.setattr(thd, "polar_group", lambda: None)
    try:
        check_exchange(state, h, nlat, L, M, B, C, dev)
        torch.cuda.synchronize()
        assert epoch[0] > 0
        # the kernel numbered the barriers itself: slot 0 (arrival of rank 0) and slot 32 (the counter) hold their count
        fl = torch.as_tensor(p2p._DevPtr(flags.value, 64), device=dev).view(torch.int32).cpu()
        assert int(fl[0]) == epoch[0] and int(fl[32]) == epoch[0]
    finally:
        torch.cuda.synchronize()
        for bufs in pool.values():
            for ptrs in bufs:
                for p in ptrs:
                    lib.sfno_ipc_free(C_.c_void_p(p))
        lib.sfno_ipc_free(flags)
