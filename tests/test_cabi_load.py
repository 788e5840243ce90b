"""The C-ABI shared library builds for sm_100a, loads, and exports every symbol include/colxlip_b200.h
declares (no compute calls: this runs without a GPU)."""
import ctypes
import os
import re
import subprocess

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "colxlip_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(cxl_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from colxlip_b200 import _lib
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
        assert n in _lib.SYMBOLS, f"{n} has no ctypes prototype"
    for n in _lib.SYMBOLS:
        assert n in names, f"{n} bound in _lib.py but missing from the header"


def test_pure_functions_and_error_strings():
    from colxlip_b200 import _lib
    lib = _lib.load()
    assert lib.cxl_version() >= 100
    assert lib.cxl_argmax_elem_bytes(196) == 1 and lib.cxl_argmax_elem_bytes(576) == 2
    assert lib.cxl_img_row_stride(196) == 196 and lib.cxl_img_row_stride(49) == 56 and lib.cxl_img_row_stride(1) == 8
    assert b"no CPU fallback" in lib.cxl_strerror(-4)
    assert lib.cxl_strerror(0) == b"ok"


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_compute_entry_fails_loudly_without_device():
    from colxlip_b200 import _lib
    lib = _lib.load()
    assert lib.cxl_device_check() == -4
    buf = (ctypes.c_float * 16)()
    rc = lib.cxl_maxsim_fwd(ctypes.addressof(buf), ctypes.addressof(buf), 1, 1, 1, 1, 64, 1, 1,
                            ctypes.addressof(buf), None, None, None)
    assert rc == -4


def test_sass_is_blackwell_native():
    """tcgen05.mma -> UTCHMMA, tcgen05.ld -> LDTM, TMA -> UTMALDG (B200_PROFILING.md)."""
    from colxlip_b200 import _lib
    out = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    for mnem in ("UTCHMMA", "LDTM", "UTMALDG"):
        assert mnem in out.stdout, mnem
    assert "HGMMA" not in out.stdout
