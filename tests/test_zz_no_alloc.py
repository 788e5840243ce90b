"""cxl_infonce_stats_static: the allocation-free variant of the K3 statistics pass (column partials in a file-scope
__device__ array inside the library, 16 MiB per device).  No counterpart in the reference -- added
for callers that cannot call cudaMalloc (graph capture, serving loops).  Called through the C-ABI
and checked against cxl_infonce_stats (caller workspace) on the same inputs, twice in a row with different sizes so that
the second call runs over the first one's stale scratch.  Collected last on purpose.  `-m gpu`."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
STATIC_BYTES = 16 << 20            # CXL_INFONCE_STATIC_SCRATCH_BYTES (include/colxlip_b200.h)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _outputs(B_loc, B, dev, fill):
    return [torch.full((n,), fill, dtype=torch.float32, device=dev) for n in (B_loc, B, B, B_loc)]


def _both(lib, B_loc, B, row0, seed, dev):
    from colxlip_b200._lib import check
    g = torch.Generator(device=dev).manual_seed(seed)
    X = torch.randn(B_loc, B, device=dev, generator=g) * 0.3
    scale = torch.tensor(1 / 0.07, device=dev)
    nbytes = int(lib.cxl_infonce_workspace_bytes(B_loc, B))
    assert 0 < nbytes <= STATIC_BYTES
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    ref = _outputs(B_loc, B, dev, float("nan"))
    got = _outputs(B_loc, B, dev, float("nan"))
    check(lib.cxl_infonce_stats(_ptr(X), _ptr(scale), B_loc, B, row0, _ptr(ref[0]), _ptr(ref[1]), _ptr(ref[2]),
                                _ptr(ref[3]), _ptr(ws), nbytes, _stream()), "cxl_infonce_stats")
    before = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)
    check(lib.cxl_infonce_stats_static(_ptr(X), _ptr(scale), B_loc, B, row0, _ptr(got[0]), _ptr(got[1]), _ptr(got[2]),
                                       _ptr(got[3]), _stream()), "cxl_infonce_stats_static")
    assert (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)) == before
    torch.cuda.synchronize(dev)
    return X, scale, ref, got


def test_static_scratch_stats_match_workspace_entry_twice_in_a_row():
    from colxlip_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    # a large call first (fills 2 MiB of the array), then a smaller, ragged one that reuses its head: stale partials of
    # the first call must not leak into the second; then the first shape again
    for B_loc, B, row0, seed in ((4096, 8191, 1000, 1), (200, 777, 577, 2), (1536, 2052, 0, 3)):
        X, scale, ref, got = _both(lib, B_loc, B, row0, seed, dev)
        for r, o, name in zip(ref, got, ("row_lse", "col_max", "col_sum", "diag")):
            assert torch.isfinite(o).all(), name
            assert torch.equal(o, r), name                    # same kernels, same launch shapes: identical bits
        # and both against plain torch (the tolerance tests/test_gpu_parity.py uses for the K3 statistics)
        Z = X.double() * scale.double()
        torch.testing.assert_close(got[0].double(), torch.logsumexp(Z, dim=1), rtol=1e-5, atol=1e-5)
        col_lse = got[1].double() + got[2].double().log()
        torch.testing.assert_close(col_lse, torch.logsumexp(Z, dim=0), rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(got[3].double(), Z[torch.arange(B_loc), row0 + torch.arange(B_loc, device=dev)],
                                   rtol=1e-6, atol=1e-6)


def test_static_scratch_rejects_oversized_and_bad_arguments():
    from colxlip_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    B_loc, B = 4096, 32768                                    # 2 * 32 * 32768 * 4 = 8 MiB fits; 16384 x 32768 does not
    assert int(lib.cxl_infonce_workspace_bytes(16384, 32768)) > STATIC_BYTES
    sentinel = _outputs(8, 8, dev, 7.0)
    dummy = torch.zeros(8, device=dev)
    rc = lib.cxl_infonce_stats_static(_ptr(dummy), _ptr(dummy), 16384, 32768, 0, _ptr(sentinel[0]), _ptr(sentinel[1]),
                                      _ptr(sentinel[2]), _ptr(sentinel[3]), _stream())
    assert rc == -6 and b"static scratch" in lib.cxl_strerror(rc)     # CXL_ERR_STATIC_SCRATCH, nothing launched
    with pytest.raises(_lib.CxlError, match="static scratch"):
        _lib.check(rc, "cxl_infonce_stats_static")
    rc = lib.cxl_infonce_stats_static(_ptr(dummy), _ptr(dummy), 8, 8, 0, _ptr(sentinel[0]), None, None,
                                      _ptr(sentinel[3]), _stream())
    assert rc == -1                                           # col_max / col_sum are required here
    torch.cuda.synchronize(dev)
    assert all(bool((t == 7.0).all()) for t in sentinel)
    assert int(lib.cxl_infonce_workspace_bytes(B_loc, B)) <= STATIC_BYTES
