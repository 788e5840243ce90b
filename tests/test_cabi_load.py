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
    assert b"workspace" in lib.cxl_strerror(-5)
    assert lib.cxl_kernel_launches() >= 0            # process-wide launch counter (0 until a kernel runs)


def test_backward_workspace_sizing(monkeypatch):
    """cxl_maxsim_bwd_workspace_bytes: scratch for the dImg row lists = images-per-chunk x super-tiles x block."""
    from colxlip_b200 import _lib
    lib = _lib.load()
    BF16, F32 = 1, 0
    # cfg-2 geometry: 2523136 rows -> 4928 super-tiles of 512 rows; block = 224 header words + 512 entries
    per_img = 4928 * (224 * 4 + 512 * 4)
    n = lib.cxl_maxsim_bwd_workspace_bytes(32768, 32768, 77, 196, 512, BF16)
    assert n % per_img == 0 and n <= 4608 << 20
    imgs = n // per_img
    assert imgs % (2 * 148) == 0 and imgs >= 296     # whole waves of the gather kernel (2 images per CTA, 148 SMs)
    # everything fits: one chunk
    assert lib.cxl_maxsim_bwd_workspace_bytes(64, 10, 77, 196, 512, BF16) == 10 * 10 * (224 * 4 + 2048)
    # shapes served by the generic kernels need none
    assert lib.cxl_maxsim_bwd_workspace_bytes(64, 10, 77, 196, 512, F32) == 0
    assert lib.cxl_maxsim_bwd_workspace_bytes(64, 10, 77, 196, 96, BF16) == 0
    assert lib.cxl_maxsim_bwd_workspace_bytes(64, 10, 4, 196, 512, BF16) == 0
    # P > 200 / P > 300 variants: wider headers (300 -> 320 words, 19*4*8 = 608 words)
    assert lib.cxl_maxsim_bwd_workspace_bytes(64, 3, 77, 300, 128, BF16) == 3 * 10 * (320 * 4 + 2048)
    monkeypatch.setenv("CXL_DIMG_V", "3")
    assert lib.cxl_maxsim_bwd_workspace_bytes(64, 3, 77, 576, 128, BF16) == 3 * 10 * (608 * 4 + 2048)
    # P > 304 defaults to the v4 gather (staggered walks, step tables): 64*77 = 4928 rows, one image per CTA, sub-tiles of
    # 768 rows (ring of 2 + two worst-case tables in 227 KB) -> 7 walks per image, each block = header (19 warps x
    # 12 u16 = 464 B, rounded to 16) + 16 B per row + one padding step per phase (19*8) + the two-step end sentinel
    monkeypatch.delenv("CXL_DIMG_V")
    blk = 464 + 16 * 768 + 16 * 19 * 8 + 32
    assert lib.cxl_maxsim_bwd_workspace_bytes(64, 3, 77, 576, 128, BF16) == 3 * 7 * (blk + 2) + 256
    monkeypatch.setenv("CXL_DIMG_V", "4")            # the v4 gather can be forced for P <= 304 too
    assert lib.cxl_maxsim_bwd_workspace_bytes(64, 10, 77, 196, 512, BF16) not in (0, 10 * 10 * (224 * 4 + 2048))
    monkeypatch.delenv("CXL_DIMG_V")
    monkeypatch.setenv("CXL_DIMG_SCRATCH_MB", "1")
    small = lib.cxl_maxsim_bwd_workspace_bytes(120, 23, 77, 196, 512, BF16)
    assert small <= 1 << 20 and small % (19 * (224 * 4 + 2048)) == 0 and small // (19 * 2944) == 18


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


def test_header_is_plain_c_and_a_c_program_links_the_library(tmp_path):
    """include/colxlip_b200.h compiles as C99 and examples/cabi_probe.c (no device needed) links and runs against
    libcolxlip_b200.so -- the boundary carries no C++ or torch types."""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("gcc not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    from colxlip_b200 import _lib
    _lib.load()
    hdr = os.path.join(root, "include", "colxlip_b200.h")
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c", hdr],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    exe = str(tmp_path / "cabi_probe")
    libdir = os.path.dirname(_lib.LIB_PATH)
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-I" + os.path.join(root, "include"),
                        os.path.join(root, "examples", "cabi_probe.c"), "-o", exe, "-L" + libdir, "-lcolxlip_b200",
                        "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "CABI_PROBE OK" in r.stdout, r.stdout + r.stderr
