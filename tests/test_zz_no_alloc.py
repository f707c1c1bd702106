"""The workspace-free weight-gradient entry point, r2d2_linear_wgrad_static.

Who needs it: users of the C ABI that are not PyTorch programs and cannot come up with a scratch
buffer -- a C/C++ actor or serving host that was handed exactly its input and output buffers, a
process whose HBM was fully budgeted at start-up, a latency-critical loop that must stay away
from cudaMalloc (it synchronises the device).  For them the split-K partials live in a 32 MiB
array inside the library.

What is checked, through the C ABI: two calls in a row (different shapes and data, so the
second one finds the first one's leftovers in the shared array) each reproduce what
r2d2_linear_wgrad computes with caller scratch, within the tolerance the other GPU tests use;
a problem too large for the built-in array is refused with a message naming both sizes.
This file sorts last on purpose: a run that stops at the first failure has already run
everything else.  It passes on its own in a fresh process and after the rest of the suite.
"""
import pytest
import torch

from util import rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _wgrad_both(M, N, K, seed):
    from pytorch_r2d2_b200 import _lib
    g = torch.Generator().manual_seed(seed)
    dy = torch.randn(M, N, generator=g).cuda()
    x = torch.randn(M, K, generator=g).cuda()
    nbytes = int(_lib.load().r2d2_linear_wgrad_scratch_bytes(M, N, K))
    scratch = torch.empty(max(nbytes, 16), dtype=torch.uint8, device="cuda")
    dw_ref, db_ref = torch.empty(N, K, device="cuda"), torch.empty(N, device="cuda")
    _lib.call("r2d2_linear_wgrad", _lib.ptr(dy, "f32"), _lib.ptr(x, "f32"), _lib.ptr(dw_ref, "f32"),
              _lib.ptr(db_ref, "f32"), M, N, K, _lib.ptr(scratch), _lib.stream_ptr())
    dw, db = torch.full((N, K), float("nan"), device="cuda"), torch.full((N,), float("nan"), device="cuda")
    _lib.call("r2d2_linear_wgrad_static", _lib.ptr(dy, "f32"), _lib.ptr(x, "f32"), _lib.ptr(dw, "f32"),
              _lib.ptr(db, "f32"), M, N, K, _lib.stream_ptr())
    torch.cuda.synchronize()
    return nbytes, (dw, db), (dw_ref, db_ref), (dy, x)


def test_static_scratch_wgrad_two_calls_in_a_row(cuda_device):
    # first call: the dueling-head layer of the BASELINE config (10 splits, 21 MB of partials)
    nbytes, (dw, db), (dw_ref, db_ref), (dy, x) = _wgrad_both(2560, 1024, 512, seed=1)
    assert nbytes > 0, "this shape is meant to exercise the scratch array"
    assert rel_err(dw, dw_ref) < TOL and rel_err(db, db_ref) < TOL
    assert rel_err(dw, dy.double().T @ x.double()) < TOL
    # second call, right behind it: smaller problem, other data; stale partials must not leak
    nbytes2, (dw2, db2), (dw2_ref, db2_ref), (dy2, x2) = _wgrad_both(640, 2048, 256, seed=2)
    assert nbytes2 > 0
    assert rel_err(dw2, dw2_ref) < TOL and rel_err(db2, db2_ref) < TOL
    assert rel_err(dw2, dy2.double().T @ x2.double()) < TOL
    # a shape that needs no scratch at all takes the same entry point
    _, (dw3, db3), (dw3_ref, db3_ref), _ = _wgrad_both(96, 70, 33, seed=3)
    assert rel_err(dw3, dw3_ref) < TOL and rel_err(db3, db3_ref) < TOL


def test_static_scratch_wgrad_refuses_oversize(cuda_device):
    from pytorch_r2d2_b200 import _lib
    M, N, K = 512, 2048, 2304            # 2 splits x 2048 x 2304 floats = 37.7 MB > 32 MiB
    assert int(_lib.load().r2d2_linear_wgrad_scratch_bytes(M, N, K)) > (32 << 20)
    dy, x = torch.zeros(M, N, device="cuda"), torch.zeros(M, K, device="cuda")
    dw = torch.zeros(N, K, device="cuda")
    with pytest.raises(RuntimeError, match=r"needs \d+ bytes.*holds 33554432"):
        _lib.call("r2d2_linear_wgrad_static", _lib.ptr(dy, "f32"), _lib.ptr(x, "f32"), _lib.ptr(dw, "f32"),
                  None, M, N, K, _lib.stream_ptr())
