"""ctypes binding of libcolxlip_b200.so (the C-ABI declared in include/colxlip_b200.h).

There is no CPU fallback: if the library cannot be loaded (and cannot be built because nvcc is
absent) every compute entry raises.  The library is built in-tree so it travels with the repo.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_longlong, c_size_t, c_ulonglong, c_void_p
from typing import Optional

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libcolxlip_b200.so")

CXL_DTYPE_F32 = 0
CXL_DTYPE_BF16 = 1
CXL_DTYPE_F16 = 2
CXL_MEAN_MASKED = 0
CXL_MEAN_PLAIN = 1
CXL_MEAN_ROUND16 = 0x10   # OR into mean_mode: row maxima rounded to the 16-bit operand dtype (autocast semantics)

_lib: Optional[ctypes.CDLL] = None

# every symbol include/colxlip_b200.h declares: (restype, argtypes)
_FP = ctypes.POINTER(c_float)
SYMBOLS = {
    "cxl_strerror": (c_char_p, [c_int]),
    "cxl_kernel_launches": (c_ulonglong, []),
    "cxl_last_cuda_error": (c_char_p, []),
    "cxl_version": (c_int, []),
    "cxl_device_check": (c_int, []),
    "cxl_argmax_elem_bytes": (c_int, [c_int]),
    "cxl_argmax_ld": (c_longlong, [c_int, c_int]),
    "cxl_img_row_stride": (c_int, [c_int]),
    "cxl_pack_bf16": (c_int, [c_void_p, c_int, c_longlong, c_longlong, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "cxl_maxsim_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                               c_void_p, c_void_p, c_void_p, c_void_p]),
    "cxl_maxsim_fwd_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "cxl_maxsim_fwd_ws": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_size_t, c_void_p]),
    "cxl_pair_weights": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_longlong, c_void_p, c_void_p]),
    "cxl_maxsim_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                               c_int, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "cxl_maxsim_bwd_ws": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_size_t, c_void_p]),
    "cxl_maxsim_bwd_dtxt_rows": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                         c_int, c_int, c_int, c_void_p, c_int, c_void_p]),
    "cxl_maxsim_bwd_shard": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                     c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_size_t,
                                     c_void_p]),
    "cxl_maxsim_bwd_dtxt_rows_shard": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                               c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "cxl_pack_operand": (c_int, [c_void_p, c_int, c_longlong, c_longlong, c_int, c_int, c_int, c_int, c_void_p, c_int,
                                 c_void_p]),
    "cxl_maxsim_bwd_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "cxl_maxsim_bwd_generic": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                       c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "cxl_infonce_workspace_bytes": (c_size_t, [c_int, c_int]),
    "cxl_infonce_stats": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_size_t, c_void_p]),
    "cxl_infonce_stats_static": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_void_p]),
    "cxl_infonce_grad": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                 c_void_p, c_void_p, c_void_p]),
    "cxl_infonce_grad_workspace_bytes": (c_size_t, [c_int, c_int]),
    "cxl_infonce_grad_ws": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "cxl_l2norm_fwd": (c_int, [c_void_p, c_int, c_longlong, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "cxl_l2norm_bwd": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_longlong, c_int, c_void_p, c_int, c_void_p]),
    "cxl_gemm_nt": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_longlong, c_void_p]),
    "cxl_cast_transpose": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "cxl_head_ln_in": (c_int, [c_void_p, c_int, c_longlong, c_int, c_void_p, c_void_p, c_float, c_void_p, c_int, c_void_p,
                               c_int, c_void_p]),
    "cxl_head_out": (c_int, [c_void_p, c_longlong, c_int, c_void_p, c_void_p, c_void_p, c_float, c_int, c_void_p, c_int,
                             c_void_p]),
    "cxl_rank_of_target": (c_int, [c_void_p, c_longlong, c_longlong, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "cxl_maxsim_fwd_tc_ex": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                     c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "cxl_debug_tile": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
}


class CxlError(RuntimeError):
    pass


def load(build_if_missing: bool = True) -> ctypes.CDLL:
    """Load (building first if needed and possible) the CUDA library.  Raises if unavailable."""
    global _lib
    if _lib is not None:
        return _lib
    if build_if_missing:
        from . import build as _build
        try:
            if _build.needs_build():
                _build.build_library()
        except Exception as e:  # nvcc missing on the box: use the shipped .so if there is one
            if not os.path.exists(LIB_PATH):
                raise CxlError(f"libcolxlip_b200.so is missing and could not be built: {e}") from e
    if not os.path.exists(LIB_PATH):
        raise CxlError("libcolxlip_b200.so not found; run `python -m colxlip_b200.build` (no CPU fallback exists)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc == 0:
        return
    lib = load()
    msg = lib.cxl_strerror(rc).decode()
    extra = lib.cxl_last_cuda_error().decode() if rc == -3 else ""
    raise CxlError(f"{what}: {msg} {extra}".strip())
