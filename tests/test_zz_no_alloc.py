"""`sfno_dhconv_fwd_static` - the allocation-free variant of the dhconv forward contraction (include/makani_b200.h) -
against the existing entry point `sfno_dhconv_fwd`, through the C-ABI.

The variant keeps the tile list and the tile scheduler's counters in a file-scope device array instead of cudaMalloc'ed
memory (makani_b200/csrc/gemm_host.cu); the array is shared by all calls and keeps the last call's contents, so the test
makes several calls in a row with DIFFERENT shapes (a larger one first, so that stale tiles of the previous call lie
behind the new list) and twice with the same shape.  The two entry points run the same GEMM kernel over the same tiles,
so the outputs - whole buffers, including the region m > l that neither call writes (pre-filled with a sentinel) - must
be bit-identical; that is inside every tolerance used elsewhere in tests/ (TF32: 1e-3 relative L2).

The file is named so that pytest collects it last; it passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env(built_lib):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from makani_b200 import _ext, ops
    return _ext, ops, _ext.load()


def _call(fn, _ext, x_sp, wp, out, B, Cin, Cout, L, M, l_off, m_off, name):
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = fn(C.c_void_p(x_sp.data_ptr()), C.c_void_p(wp.data_ptr()), C.c_void_p(out.data_ptr()), B, Cin, Cout, L, M,
            l_off, m_off, 1, stream)
    _ext.check(rc, name)


# (B, Cin, Cout, L, M, l_offset, m_offset): larger first, then smaller (stale tiles behind the list), then a sharded
# layout with offsets, then the first shape again
CASES = [
    (2, 40, 136, 48, 49, 0, 0),
    (1, 16, 16, 20, 21, 0, 0),
    (2, 24, 20, 16, 12, 16, 5),
    (2, 40, 136, 48, 49, 0, 0),
]


def test_static_variant_equals_the_allocating_entry_point_on_calls_in_a_row(env):
    _ext, ops, lib = env
    dev = torch.device("cuda:0")
    torch.manual_seed(11)
    for it, (B, Cin, Cout, L, M, lo, mo) in enumerate(CASES):
        x = torch.randn(B, Cin, L, M, dtype=torch.complex64, device=dev)
        w = torch.randn(Cin, Cout, L, dtype=torch.complex64, device=dev)
        x_sp = ops.c64_to_sp(x, lo, mo)
        wp = ops.pack_weight(w)
        n = ops.sp_elems(L, M, B, Cout)
        ref = torch.full((n,), 7.25, dtype=torch.float32, device=dev)
        out = torch.full((n,), 7.25, dtype=torch.float32, device=dev)
        _call(lib.sfno_dhconv_fwd, _ext, x_sp, wp, ref, B, Cin, Cout, L, M, lo, mo, "sfno_dhconv_fwd")
        for rep in range(2):                                   # two calls in a row on the same inputs
            out.fill_(7.25)
            _call(lib.sfno_dhconv_fwd_static, _ext, x_sp, wp, out, B, Cin, Cout, L, M, lo, mo, "sfno_dhconv_fwd_static")
            torch.cuda.synchronize()
            assert torch.equal(out, ref), (it, rep, float((out - ref).abs().max()))
        assert bool((ref != 7.25).any())                       # the contraction did write something


def test_static_variant_interleaved_with_the_allocating_one(env):
    """The allocating entry point uses its own counters; alternating the two on one stream must not disturb either."""
    _ext, ops, lib = env
    dev = torch.device("cuda:0")
    torch.manual_seed(12)
    B, Cin, Cout, L, M = 1, 384, 384, 240, 241                 # the named config's contraction (1 056 tiles)
    x = torch.randn(B, Cin, L, M, dtype=torch.complex64, device=dev)
    w = torch.randn(Cin, Cout, L, dtype=torch.complex64, device=dev) / Cin ** 0.5
    x_sp = ops.c64_to_sp(x, 0, 0)
    wp = ops.pack_weight(w)
    n = ops.sp_elems(L, M, B, Cout)
    outs = [torch.zeros(n, dtype=torch.float32, device=dev) for _ in range(4)]
    for i, o in enumerate(outs):
        fn, name = ((lib.sfno_dhconv_fwd, "sfno_dhconv_fwd") if i % 2 == 0 else
                    (lib.sfno_dhconv_fwd_static, "sfno_dhconv_fwd_static"))
        _call(fn, _ext, x_sp, wp, o, B, Cin, Cout, L, M, 0, 0, name)
    torch.cuda.synchronize()
    for o in outs[1:]:
        assert torch.equal(o, outs[0])
    # against the einsum of the reference (contractions.py:131-136) in fp64, on m <= l: the tests' TF32 tolerance
    y = ops.sp_to_c64(outs[1], B, Cout, L, M).to(torch.complex128)
    mask = (torch.arange(M, device=dev).view(1, M) <= torch.arange(L, device=dev).view(L, 1))
    ref = torch.einsum("bixy,iox->boxy", (x * mask).to(torch.complex128), w.to(torch.complex128))
    err = float((y * mask - ref).norm() / ref.norm())
    assert err < 1e-3, err


def test_static_variant_rejects_shapes_beyond_its_limits(env):
    _ext, ops, lib = env
    dev = torch.device("cuda:0")
    dummy = torch.zeros(1024, dtype=torch.float32, device=dev)
    p = C.c_void_p(dummy.data_ptr())
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # B = 32 of the named config needs 23 040 tiles > SFNO_STATIC_MAX_TILES; rejected before anything is launched
    assert lib.sfno_dhconv_fwd_static(p, p, p, 32, 384, 384, 240, 241, 0, 0, 1, s) == 1
    assert "SFNO_STATIC_MAX_TILES" in lib.sfno_last_error().decode()
    assert lib.sfno_dhconv_fwd_static(p, p, p, 1, 8, 8, 1025, 8, 0, 0, 1, s) == 1
    assert "SFNO_STATIC_MAX_L" in lib.sfno_last_error().decode()
    torch.cuda.synchronize()
