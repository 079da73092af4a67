"""ctypes binding of libkanvit_b200.so (the C ABI declared in include/kanvit_b200.h).

There is no CPU fallback: importing the ops on a machine without the built library, or
calling them with non-CUDA tensors, raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libkanvit_b200.so")

# kv_family
KV_LINEAR, KV_EFFKAN, KV_CHEBY, KV_FASTKAN, KV_SINE, KV_FOURIER = range(6)
KV_MAX_M = 64
KV_PTRS_PER_GROUP = 8
KV_FLAG_RELU = 1
KV_FLAG_NO_LAYERNORM = 2
KV_FLAG_NO_BASE = 4
KV_FLAG_BF16_TC = 8
KV_FLAG_UNIFORM_GRID = 16
KV_FLAG_ACCUM_GRADS = 32
KV_ERR_UNSUPPORTED = -5


class KanDesc(C.Structure):
    """kv_kan_desc_t"""
    _fields_ = [
        ("family", C.c_int32), ("T", C.c_int32), ("in_features", C.c_int32),
        ("out_features", C.c_int32), ("n_groups", C.c_int32), ("x_share", C.c_int32),
        ("M", C.c_int32), ("grid_size", C.c_int32), ("spline_order", C.c_int32),
        ("ldx", C.c_int32), ("ldy", C.c_int32),
        ("rows_in", C.c_int32), ("rows_out", C.c_int32), ("row_shift", C.c_int32),
        ("flags", C.c_uint32), ("rbf_denom", C.c_float),
        ("fparam_stride", C.c_int32), ("reserved", C.c_int32),
    ]


class AttnGeneral(C.Structure):
    """kv_attn_general_t of include/kanvit_b200.h"""
    _fields_ = [(n, C.c_int32) for n in ("B", "H", "Nq", "Nk", "d_head", "ldq", "ldk", "ldv", "ldo", "causal")] + \
               [(n, C.c_int64) for n in ("mask_stride_b", "mask_stride_h", "mask_stride_q")]


_P = C.c_void_p
_I = C.c_int
_DP = C.POINTER(KanDesc)

# name -> argtypes; every function returns int (kv_status) unless noted
_SIGNATURES = {
    "kv_kan_fparam_count": [_DP],
    "kv_kan_pack": [_DP, _P, _P, _P, _P, _P, _P, _P],
    "kv_kan_tc_bytes": [_DP, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)],
    "kv_kan_dy_image": [_DP, _P, _P, _P, _P],
    "kv_row_stats": [_P, _I, _I, _I, _I, C.c_float, _P, _P, _P],
    "kv_kan_fwd": [_DP] + [_P] * 11,
    "kv_kan_bwd_input": [_DP] + [_P] * 13,
    "kv_kan_bwd_weight": [_DP] + [_P] * 10,
    "kv_fastkan_ln_bwd": [_DP] + [_P] * 8,
    "kv_kan_unpack_grads": [_DP] + [_P] * 6,
    "kv_attn_fwd": [_P, _P, _P, _P, _I, _I, _I, _I, _P],
    "kv_attn_bwd": [_P, _P, _P, _P, _I, _I, _I, _I, _P],
    "kv_attn_fwd_bf16": [_P, _P, _P, _P, _I, _I, _I, _I, _P],
    "kv_attn_bwd_bf16": [_P, _P, _P, _P, _I, _I, _I, _I, _P],
    "kv_kan_fwd_attn_image": [_DP] + [_P] * 7 + [_I, _P],
    "kv_kan_fwd_noalloc": [_DP] + [_P] * 8,
    "kv_kan_dy_image_geom": [_DP, C.POINTER(C.c_int), C.POINTER(C.c_int)],
    "kv_attn_image_fwd": [_P, _P, _P, _P, _I, _I, _I, _I, _P],
    "kv_attn_image_bwd": [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _P],
    "kv_attn_general_fwd": [C.POINTER(AttnGeneral)] + [_P] * 7,
    "kv_attn_general_bwd": [C.POINTER(AttnGeneral)] + [_P] * 11,
    "kv_layernorm_fwd": [_P, _P, _P, _P, _P, _P, _I, _I, _I, C.c_float, _P],
    "kv_layernorm_bwd": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, C.c_float, _P],
    "kv_patchify": [_P, _P, _I, _I, _I, _I, _I, _P],
    "kv_patchify_bwd": [_P, _P, _I, _I, _I, _I, _I, _P],
    "kv_cls_rows_fwd": [_P, _P, _P, _I, _I, _I, _P],
    "kv_cls_rows_bwd": [_P, _P, _I, _I, _I, _P],
    "kv_gather_rows": [_P, _P, _I, _I, _I, _P],
    "kv_scatter_rows": [_P, _P, _I, _I, _I, _P],
    "kv_cross_entropy_workspace_floats": [],
    "kv_cross_entropy": [_P, _P, _I, _I, C.c_longlong, _P, _P, _P, _P, _P, _P, _P],
    "kv_image_augment": [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P],
    "kv_adam_step": [_P, _P, _P, _P, C.c_longlong, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, _I,
                     _P, _P],
}
EXPORTS = ["kv_version", "kv_status_string", "kv_last_cuda_error", "kv_launch_count", "kv_attn_image_bytes"] + \
    list(_SIGNATURES)

_lib = None


class KanVitError(RuntimeError):
    pass


def load():
    """Load the shared library once; raise if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise KanVitError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C kan_vit_b200/csrc` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    lib.kv_version.restype = C.c_int
    lib.kv_version.argtypes = []
    lib.kv_status_string.restype = C.c_char_p
    lib.kv_status_string.argtypes = [C.c_int]
    lib.kv_last_cuda_error.restype = C.c_char_p
    lib.kv_last_cuda_error.argtypes = []
    lib.kv_launch_count.restype = C.c_ulonglong
    lib.kv_launch_count.argtypes = []
    for name, args in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = C.c_int
        fn.argtypes = args
    lib.kv_attn_image_bytes.restype = C.c_size_t
    lib.kv_attn_image_bytes.argtypes = [C.c_int] * 4
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc == 0:
        return
    lib = load()
    msg = lib.kv_status_string(rc).decode()
    if rc == -6:
        msg += ": " + lib.kv_last_cuda_error().decode()
    raise KanVitError(f"{what} failed: {msg} ({rc})")


def ptr(t):
    """Device pointer of a tensor or None."""
    return None if t is None else t.data_ptr()
