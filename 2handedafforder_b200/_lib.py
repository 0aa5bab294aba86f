"""ctypes binding of libb200sam.so (C ABI declared in include/b200sam.h).

There is no CPU or PyTorch fallback: if the shared library is missing this
module raises at first use, and every non-zero status becomes a RuntimeError
carrying `b200sam_last_error`.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libb200sam.so")

F32, BF16, F16 = 0, 1, 2
VIT_B, VIT_L, VIT_H = 0, 1, 2

# every symbol include/b200sam.h, include/b200sam_train.h and include/b200sam_debug.h declare: (name, restype, argtypes)
_P, _I, _L, _F = C.c_void_p, C.c_int, C.c_int64, C.c_float
SYMBOLS = [
    ("b200sam_create", _I, [C.POINTER(_P), _I, _I, _I, _I]),
    ("b200sam_create_custom", _I, [C.POINTER(_P), _I, _I, _I, C.POINTER(_I), _I, _I, _I, _I]),
    ("b200sam_destroy", None, [_P]),
    ("b200sam_last_error", C.c_char_p, [_P]),
    ("b200sam_load_weight", _I, [_P, C.c_char_p, _P, _I, C.POINTER(_L), _I]),
    ("b200sam_finalize_weights", _I, [_P]),
    ("b200sam_encode", _I, [_P, _P, _I, _I, _P, _I, _P]),
    ("b200sam_decode", _I, [_P, _P, _I, _I, _P, C.POINTER(_I), _P, _P, _P, _P, _P, _P]),
    ("b200sam_postprocess", _I, [_P, _P, _I, _I, _I, _I, _I, _P, _P, _F, _P]),
    ("b200sam_postprocess_bits", _I, [_P, _P, _I, _I, _I, _I, _I, _P, _F, _P]),
    ("b200sam_postprocess_pair", _I, [_P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _F, _P, _P, _P, _P, _P, _P]),
    ("b200sam_checksum64", _I, [_P, _P, _I, _L, _L, _P, _P]),
    ("b200sam_dense_pe", _I, [_P, _P, _P]),
    ("b200sam_preprocess", _I, [_P, _P, _I, _I, _I, _P, _I, C.POINTER(_I), _P]),
    ("b200sam_seg_project", _I, [_P, _P, _I, _I, _I, _I, _P, _I, _L, _P, _P, _P, _P, _I, _P, _I, _P, _P]),
    ("b200sam_seg_project_noalloc", _I, [_P, _P, _I, _I, _I, _I, _P, _I, _L, _P, _P, _P, _P, _I, _P, _I, _P, _P]),
    ("b200sam_mask_metrics", _I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _P, _P, _P]),
    ("b200sam_launch_count", _L, [_P]),
    ("b200sam_set_debug_simt_gemm", _I, [_P, _I]),
    ("b200sam_profile_enable", _I, [_P, _I]),
    ("b200sam_profile_read", _I, [_P, C.POINTER(_F), C.POINTER(_L), _I]),
    ("b200sam_debug_stop_after_block", _I, [_P, _I]),
    ("b200sam_debug_read_tokens", _I, [_P, _P, _I, _P]),
    ("b200sam_op_gemm", _I, [_I, _I, _I, _P, _I, _P, _I, _I, _I, _I, _P, _P, _I, _I, _P, _P, _F, _P, _I, _I, _I, _P]),
    ("b200sam_op_gemm_fold_producer", _I, [_I, _P, _I, _P, _I, _I, _I, _I, _P, _P, _I, _I, _P, _I, _P, _P, _P]),
    ("b200sam_op_gemm_fold_consumer", _I, [_I, _I, _P, _I, _P, _P, _P, _P, _F, _P, _I, _I, _I, _P, _P, _P, _I, _P]),
    ("b200sam_op_layernorm", _I, [_P, _I, _I, _P, _P, _F, _P, _I, _P]),
    ("b200sam_op_attention", _I, [_P, _I, _I, _I, _I, _P, _P, _P, _P]),
    # include/b200sam_train.h (fp32 forward / backward primitives of the decoder training graph)
    ("b200sam_train_gemm", _I, [_P, _P, C.POINTER(_L), _P, C.POINTER(_L), _P, C.POINTER(_L), _P, _I, _I, _I, _I, _I, _I,
                                _F, _F, _I, _P]),
    ("b200sam_train_gemm_tc", _I, [_P, _P, _L, _P, _L, _P, _L, _P, _I, _I, _I, _I, _P]),
    ("b200sam_train_transpose", _I, [_P, _P, _L, _I, _P, _P]),
    ("b200sam_train_layernorm_fwd", _I, [_P, _P, _P, _P, _L, _I, _F, _P, _P, _P, _P]),
    ("b200sam_train_layernorm_bwd", _I, [_P, _P, _P, _P, _P, _P, _L, _I, _P, _P, _P, _P]),
    ("b200sam_train_softmax_fwd", _I, [_P, _P, _L, _I, _P, _P]),
    ("b200sam_train_softmax_bwd", _I, [_P, _P, _P, _L, _I, _P, _P]),
    ("b200sam_train_act_fwd", _I, [_P, _I, _P, _L, _P, _P]),
    ("b200sam_train_act_bwd", _I, [_P, _I, _P, _P, _L, _P, _P]),
    ("b200sam_train_add_bcast", _I, [_P, _P, _P, _L, _I, _L, _P, _P]),
    ("b200sam_train_colsum", _I, [_P, _P, _L, _I, _P, _P]),
    ("b200sam_train_pixel_shuffle", _I, [_P, _P, _L, _I, _I, _I, _I, _P, _P]),
    ("b200sam_train_src_tokens", _I, [_P, _P, _I, _I, _P, _P, _I, _P, _P]),
    ("b200sam_train_resize_fwd", _I, [_P, _P, _L, _I, _I, _L, _L, _I, _I, _P, _P]),
    ("b200sam_train_resize_bwd", _I, [_P, _P, _L, _I, _I, _L, _L, _I, _I, _P, _P]),
    ("b200sam_train_mask_loss_fwd", _I, [_P, _P, _P, _P, _I, _L, _P, _P]),
    ("b200sam_train_mask_loss_bwd", _I, [_P, _P, _P, _P, _P, _P, _P, _I, _L, _F, _F, _P, _P]),
    # include/b200sam_debug.h (test / measurement hooks, not part of the drop-in surface)
    ("b200sam_debug_set_variant", _I, [_I, _I]),
    ("b200sam_debug_set_trace", _I, [_P]),
    ("b200sam_debug_attention_bench", _I, [_P, _I, _I, _I, _I, _P, _P, _P, _I, C.POINTER(_F), _P]),
]

_lib: Optional[C.CDLL] = None


def lib_path() -> str:
    return os.environ.get("B200SAM_LIB", LIB_PATH)


def load() -> C.CDLL:
    """dlopen the library and bind every declared symbol (fails loudly if absent)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(libb200sam has no CPU/PyTorch fallback)")
    lib = C.CDLL(path)
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)   # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error(handle) -> str:
    msg = load().b200sam_last_error(handle)
    return msg.decode() if msg else ""


def check(rc: int, handle=None, what: str = "") -> None:
    if rc != 0:
        detail = last_error(handle) if handle else ""
        raise RuntimeError(f"libb200sam {what} failed (status {rc}): {detail}")
