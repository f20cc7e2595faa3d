"""ctypes binding of the C-ABI library ``libmakani_b200.so`` (include/makani_b200.h).

There is no fallback: if the library is missing or a call fails, this raises.  Build it in-tree with
``python -c "import __graft_entry__ as g; g.build()"`` (or ``make -C makani_b200/csrc``).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmakani_b200.so")

F32, BF16, F16 = 0, 1, 2

_vp, _i, _ll = C.c_void_p, C.c_int, C.c_longlong

# name -> (restype, argtypes); mirrors include/makani_b200.h line by line
_PROTOTYPES = {
    "sfno_abi_version": (_i, []),
    "sfno_last_error": (C.c_char_p, []),
    "sfno_launch_count": (_ll, []),
    "sfno_plan_create": (_i, [_i, _i, _i, _i, _i, _vp, C.POINTER(_vp)]),
    "sfno_plan_destroy": (_i, [_vp]),
    "sfno_xm_elems": (_ll, [_i, _i, _i, _i]),
    "sfno_sp_elems": (_ll, [_i, _i, _i, _i]),
    "sfno_wp_elems": (_ll, [_i, _i, _i]),
    "sfno_dft_fwd": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "sfno_dft_inv": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "sfno_legendre_analysis": (_i, [_vp, _vp, _i, _i, _vp, _i, _vp]),
    "sfno_legendre_synthesis": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "sfno_dft_fwd_win": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "sfno_dft_inv_win": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "sfno_legendre_analysis_win": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _i, _i, _vp]),
    "sfno_legendre_synthesis_win": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "sfno_dhconv_pack_weight": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "sfno_dhconv_pack_weight_ex": (_i, [_vp, _i, _i, _i, _vp, _i, _vp]),
    "sfno_dhconv_unpack_wgrad": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "sfno_dhconv_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "sfno_dhconv_fwd_static": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "sfno_dhconv_bwd_x": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "sfno_dhconv_bwd_w": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "sfno_diag_contract": (_i, [_i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "sfno_sp_to_c64": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "sfno_c64_to_sp": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "sfno_xm_to_c64": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "sfno_c64_to_xm": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "sfno_add_bias": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp]),
    "sfno_dhconv_unpack_wgrad_gather": (_i, [_vp, _i, _ll, _i, _i, _i, _vp, _vp]),
    "sfno_peer_copy": (_i, [_vp, _vp, C.c_size_t, _vp]),
    "sfno_slices_average": (_i, [_vp, _vp, _i, _ll, C.c_float, _vp]),
    "sfno_add_inplace": (_i, [_vp, _vp, _ll, _vp]),
    "sfno_dft_fwd_band": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "sfno_dft_inv_band": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "sfno_xm_mirror_conj": (_i, [_vp, _vp, _i, _i, _ll, _vp, _i, _vp]),
    "sfno_sp_degree_power": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "sfno_sp_degree_power_bwd": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "sfno_xm_add_dc": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "sfno_workspace_bytes": (C.c_size_t, [_vp, _vp, _i, _i, _i]),
    "sfno_sht_fwd": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, C.c_size_t, _vp]),
    "sfno_sht_inv": (_i, [_vp, _vp, _i, _i, _vp, _i, _vp, C.c_size_t, _vp]),
    "sfno_block_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, C.c_size_t, _vp]),
    "sfno_ipc_alloc": (_i, [C.c_size_t, C.POINTER(_vp), C.c_char_p]),
    "sfno_ipc_open": (_i, [C.c_char_p, C.POINTER(_vp)]),
    "sfno_ipc_close": (_i, [_vp]),
    "sfno_ipc_free": (_i, [_vp]),
    "sfno_rows_scatter": (_i, [_vp, _vp, _i, _vp]),
    "sfno_rows_copy_ce": (_i, [_vp, _vp, _i, _vp]),
    "sfno_peer_barrier": (_i, [_vp, _i, _i, _i, _vp]),
    "sfno_dft_peer_supported": (_i, [_i, _i]),
    "sfno_dft_fwd_peer": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "sfno_dft_inv_peer": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp]),
    "sfno_instnorm_stats": (_i, [_vp, _i, _i, _ll, C.c_float, _vp, _vp, _vp, _vp]),
    "sfno_instnorm_apply": (_i, [_vp, _vp, _i, _i, _i, _ll, _vp, _vp, _vp, _i, _vp, _vp]),
    "sfno_instnorm_bwd_stats": (_i, [_vp, _vp, _i, _i, _i, _ll, _vp, _vp, _vp, _i, _vp, _vp]),
    "sfno_instnorm_bwd_apply": (_i, [_vp, _vp, _i, _i, _i, _ll, _ll, _vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "sfno_bias_act": (_i, [_vp, _i, _i, _i, _ll, _vp, _i, _vp, _vp]),
    "sfno_bias_act_bwd": (_i, [_vp, _vp, _i, _i, _i, _ll, _vp, _i, _vp, _vp, _vp]),
}


class CopyDesc(C.Structure):
    """sfno_copy_desc (include/makani_b200.h)."""
    _fields_ = [("dst", _vp), ("src_off", _ll), ("dst_off", _ll),
                ("n0", _i), ("n1", _i), ("n2", _i), ("len", _i),
                ("ss0", _ll), ("ss1", _ll), ("ss2", _ll), ("ds0", _ll), ("ds1", _ll), ("ds2", _ll),
                ("tri_l0", _i)]



OP_SHT_FWD, OP_SHT_INV, OP_BLOCK_FWD = 0, 1, 2
INSTNORM_SEGMENTS = 8
ACT_NONE, ACT_GELU = 0, 1
M_OFFSET_DENSE, DENSE_L_OFFSET = -1, 1 << 20          # SFNO_M_OFFSET_DENSE / SFNO_DENSE_L_OFFSET
ABI_VERSION = 3

EXPORTED_SYMBOLS = tuple(_PROTOTYPES)

_lib = None


def load():
    """Load the shared library (once) and attach prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA extension has not been built and makani_b200 has no CPU fallback. "
            "Run `python -c \"import __graft_entry__ as g; g.build()\"` from the repository root.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _PROTOTYPES.items():
        fn = getattr(lib, name)          # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    if lib.sfno_abi_version() != ABI_VERSION:
        raise ImportError(f"{LIB_PATH}: ABI version {lib.sfno_abi_version()} != {ABI_VERSION}")
    _lib = lib
    return lib


def check(status: int, what: str):
    if status != 0:
        msg = load().sfno_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"makani_b200: {what} failed: {msg}")


def launch_count() -> int:
    return int(load().sfno_launch_count())
