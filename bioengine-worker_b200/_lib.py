"""ctypes binding of libbe_b200.so (the C-ABI declared in include/be_b200.h).

There is no fallback: if the library is missing or a call fails, a `BeError` is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbe_b200.so")

BE_OK, BE_ERR_INVALID_ARG, BE_ERR_CUDA, BE_ERR_TOO_LARGE = 0, -1, -2, -3
BE_OP_CLIP, BE_OP_SCALE_LINEAR, BE_OP_SIGMOID = 0, 1, 2
BE_EPI_NONE, BE_EPI_SIGMOID = 0, 1


class BeError(RuntimeError):
    """A be_* entry point returned a negative code."""

    def __init__(self, fn: str, code: int, msg: str):
        super().__init__(f"{fn} failed ({code}): {msg}")
        self.code = code


_i64, _i32, _f32, _f64, _ptr, _sz = C.c_int64, C.c_int, C.c_float, C.c_double, C.c_void_p, C.c_size_t

# name -> (restype, argtypes); mirrors include/be_b200.h one to one
SIGNATURES = {
    "be_version": (_i32, []),
    "be_last_error_string": (C.c_char_p, []),
    "be_meanvar_workspace_bytes": (_sz, [_i64, _i64]),
    "be_meanvar_f32": (_i32, [_ptr, _i64, _i64, _f64, _ptr, _ptr, _ptr, _ptr]),
    "be_quantile_workspace_bytes": (_sz, [_i64, _i64]),
    "be_quantile_f32": (_i32, [_ptr, _i64, _i64, _ptr, _i32, _f64, _ptr, _ptr, _ptr, _ptr]),
    "be_quantile_f32_noalloc": (_i32, [_ptr, _i64, _i64, _ptr, _i32, _f64, _ptr, _ptr, _ptr]),
    "be_normalize_f32": (_i32, [_ptr, _ptr, _i64, _i64, _ptr, _i32, _f32, _f32, _ptr]),
    "be_elementwise_f32": (_i32, [_ptr, _ptr, _i64, _i32, _f32, _f32, _ptr]),
    "be_scale_linear_seg_f32": (_i32, [_ptr, _ptr, _i64, _i64, _ptr, _ptr]),
    "be_cast_f32": (_i32, [_ptr, _i32, _ptr, _i64, _ptr]),
    "be_binarize_f32_u8": (_i32, [_ptr, _ptr, _i64, _f32, _ptr]),
    "be_tile_gather_f32": (_i32, [_ptr, _ptr] + [_i64] * 5 + [_i32] * 6 + [_ptr] * 4 + [_i32, _i32, _f32, _f32, _ptr]),
    "be_stitch_crop_f32": (_i32, [_ptr] * 4 + [_i64] * 5 + [_i32] * 6 + [_ptr] * 3 + [_i32, _f32, _ptr]),
    "be_stitch_blend_f32": (_i32, [_ptr] * 4 + [_i64] * 5 + [_i32] * 9 + [_ptr] * 4 + [_i32, _f32, _ptr]),
    "be_blend_window_f32": (_i32, [_ptr] * 3 + [_i32, _i32, _ptr, _ptr]),
    "be_ccl_workspace_bytes": (_sz, [_i64] * 4),
    "be_ccl_u8_i32": (_i32, [_ptr, _ptr] + [_i64] * 4 + [_i32, _ptr, _ptr, _ptr]),
    "be_ccl_bits_i32": (_i32, [_ptr, _ptr] + [_i64] * 4 + [_i32, _ptr, _ptr, _ptr]),
    "be_ccl_f32_i32": (_i32, [_ptr, _i32, _f32, _ptr] + [_i64] * 4 + [_i32, _ptr, _ptr, _ptr]),
    "be_label_overlap_i32": (_i32, [_ptr, _ptr, _i64, _i64, _i64, _ptr, _ptr, _i64, _ptr, _i64, _ptr, _i64, _ptr, _ptr, _ptr,
                                    _ptr]),
    "be_match_workspace_bytes": (_sz, [_i64, _i64, _i64, _i32]),
    "be_match_counts": (_i32, [_ptr, _ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i32, _ptr, _ptr, _ptr, _ptr]),
    "be_screen_rows_i64": (_i32, [_ptr, _ptr, _ptr, _i64, _i32, _ptr, _ptr, _ptr, _ptr, _ptr]),
    "be_binary_confusion_f32": (_i32, [_ptr, _ptr, _i64, _i32, _f32, _f32, _ptr, _ptr, _ptr]),
    "be_label_area_i32": (_i32, [_ptr, _i64, _i64, _i64, _ptr, _ptr, _ptr]),
    "be_keep_large_u8": (_i32, [_ptr, _ptr, _i64, _i64, _i64, _i32, _ptr, _ptr]),
    "be_binary_closing_workspace_bytes": (_sz, [_i64, _i64, _i64]),
    "be_binary_closing_u8": (_i32, [_ptr, _ptr, _i64, _i64, _i64, _ptr, _i32, _ptr, _ptr]),
    "be_edt_workspace_bytes": (_sz, [_i64, _i64, _i64]),
    "be_edt_u8_f64": (_i32, [_ptr, _ptr, _i64, _i64, _i64, _ptr, _ptr]),
    "be_edt_sq_u8_i32": (_i32, [_ptr, _ptr, _i64, _i64, _i64, _ptr, _ptr]),
    "be_peak_local_max_workspace_bytes": (_sz, [_i64, _i64]),
    "be_peak_local_max_i32": (_i32, [_ptr, _ptr, _ptr, _i64, _i64, _i32, _ptr, _ptr]),
    "be_watershed_workspace_bytes": (_sz, [_i64, _i64]),
    "be_watershed_i32": (_i32, [_ptr, _i32, _ptr, _ptr, _ptr, _ptr, _i64, _i64, _ptr, _ptr]),
    "be_label_stats_i32": (_i32, [_ptr] + [_i64] * 5 + [_ptr] * 4),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library (once) and attach prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise BeError("load", BE_ERR_INVALID_ARG,
                      f"{LIB_PATH} not found — build it with `python __graft_entry__.py` (make -C bioengine-worker_b200/csrc)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header and library out of sync
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def call(name: str, *args):
    """Call an int-returning entry point and raise BeError on a negative code."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != BE_OK:
        raise BeError(name, rc, lib.be_last_error_string().decode("utf-8", "replace"))
    return rc


def query(name: str, *args) -> int:
    return int(getattr(load(), name)(*args))
