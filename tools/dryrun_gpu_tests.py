"""Dry-run of GPU test files on a machine WITHOUT a GPU: `.cuda()` / device arguments become no-ops and the kernel
backend is the torch stand-in of tests/ref_kernels.py.  Development aid only -- it catches Python-level mistakes
(names, shapes, dtypes, control flow) in new `-m gpu` tests before GPU minutes are spent on them; it proves nothing
about the CUDA kernels.   python tools/dryrun_gpu_tests.py [tests/test_gpu_widen.py] [-k expr]"""
import sys, contextlib
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
torch.Tensor.cuda = lambda self, *a, **k: self
_to = torch.Tensor.to
def to(self, *a, **k):
    k.pop("device", None); k.pop("non_blocking", None)
    a = tuple(x for x in a if not isinstance(x, torch.device))
    return _to(self, *a, **k) if (a or k) else self
torch.Tensor.to = to
torch.nn.Module.cuda = lambda self, *a, **k: self
torch.cuda.synchronize = lambda *a, **k: None
torch.cuda.is_available = lambda: True
_tensor = torch.tensor
def tensor(*a, **k):
    k.pop("device", None)
    return _tensor(*a, **k)
torch.tensor = tensor
_autocast = torch.autocast
class FakeAutocast(contextlib.ContextDecorator):
    def __init__(self, *a, **k):
        self.en = k.get("enabled", True)
    def __enter__(self):
        import colxlip_b200.loss as L
        self.prev = L.get_precision()
        if self.en: L.set_precision("bf16")
    def __exit__(self, *e):
        import colxlip_b200.loss as L
        L.set_precision(self.prev)
torch.autocast = FakeAutocast
_flag = {"on": False}
_e, _x = FakeAutocast.__enter__, FakeAutocast.__exit__
def _enter(self):
    _flag["on"] = self.en; return _e(self)
def _exit(self, *a):
    _flag["on"] = False; return _x(self, *a)
FakeAutocast.__enter__, FakeAutocast.__exit__ = _enter, _exit
torch.is_autocast_enabled = lambda *a: _flag["on"]
torch.get_autocast_dtype = lambda *a: torch.bfloat16
_is_cuda = property(lambda self: True)
torch.Tensor.is_cuda = _is_cuda
from tests.ref_kernels import TorchRefKernels
class Dry(TorchRefKernels):
    n = 0
    def launches(self):
        Dry.n += 1
        return Dry.n
    def prepare(self, x, n_pad, prec):
        if x.dtype == torch.bfloat16 and x.is_contiguous():
            return x
        return super().prepare(x, n_pad, prec)
    def maxsim_fwd(self, txt_op, img_op, *a, **k):
        return super().maxsim_fwd(txt_op.float(), img_op.float(), *a, **k)
    def maxsim_bwd(self, txt_op, img_op, *a, **k):
        k.pop("library_scratch", None)
        return super().maxsim_bwd(txt_op.float(), img_op.float(), *a, **k)
    def maxsim_bwd_dtxt_rows(self, img_op, *a, **k):
        return super().maxsim_bwd_dtxt_rows(img_op.float(), *a, **k)
import colxlip_b200.kernels as KK
d = Dry()
KK.default_kernels = lambda: d
import colxlip_b200.loss as L, colxlip_b200.model as M, colxlip_b200.retrieval as R
L.default_kernels = M.default_kernels = R.default_kernels = lambda: d
import pytest
args = sys.argv[1:]
if not any(a.endswith(".py") for a in args):
    args = [os.path.join(ROOT, "tests", "test_gpu_widen.py")] + args
sys.exit(pytest.main(["-x", "-q", "-p", "no:cacheprovider"] + args))
