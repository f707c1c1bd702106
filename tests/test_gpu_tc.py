"""GPU: the bf16 tcgen05 GEMM path against fp64 matmul on the same bf16-rounded operands
(exactness of the tensor-core pipeline: only fp32 accumulation order differs), plus the cast /
transpose helpers.  Tolerance 1e-4 relative to max|C| on identical bf16 inputs; the end-to-end
bf16-vs-reference tolerance (2e-2, BASELINE.json) is checked in test_gpu_net_bf16."""
import pytest
import torch

from util import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (128, 128, 128), (256, 256, 512), (100, 132, 72),
                                    (5440, 2048, 1568), (2560, 1024, 512), (2560, 1568, 2048),
                                    (2048, 1568, 2560), (64, 2048, 512)])
def test_tc_linear_vs_fp64(cuda_device, M, N, K):
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(M + N + K)
    x = torch.randn(M, K, generator=g).cuda().bfloat16()
    w = (torch.randn(N, K, generator=g) * 0.1).cuda().bfloat16()
    b1, b2 = torch.randn(N, generator=g).cuda(), torch.randn(N, generator=g).cuda()
    y32 = torch.full((M, N), float("nan"), device="cuda")
    y16 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    engine.tc_linear(x, w, b1, b2, y32=y32, y16=y16)
    want = x.double() @ w.double().T + b1.double() + b2.double()
    assert rel_err(y32, want) < 1e-4, rel_err(y32, want)
    assert rel_err(y16.float(), want) < 1e-2
    y32r = torch.empty(M, N, device="cuda")
    engine.tc_linear(x, w, None, None, y32=y32r, relu=True)
    assert rel_err(y32r, torch.relu(x.double() @ w.double().T)) < 1e-4


@pytest.mark.parametrize("M,N,K", [(256, 128, 64), (5440, 2048, 1568), (2560, 1024, 512), (100, 132, 72)])
def test_tc_linear_x3_is_fp32_accurate(cuda_device, M, N, K):
    """Error-compensated bf16x3: fp32 operands split into (hi, lo); result within 1e-4 of the
    fp64 product of the ORIGINAL fp32 operands (plain bf16 would be ~4e-3)."""
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(M * 3 + N + K)
    x = torch.randn(M, K, generator=g).cuda()
    w = (torch.randn(N, K, generator=g) * 0.1).cuda()
    b = torch.randn(N, generator=g).cuda()
    xs = torch.empty(2, M, K, device="cuda", dtype=torch.bfloat16)
    ws = torch.empty(2, N, K, device="cuda", dtype=torch.bfloat16)
    engine.split_bf16(x, xs[0], xs[1])
    engine.split_bf16(w, ws[0], ws[1])
    assert torch.equal(xs[0], x.bfloat16()) and torch.equal(xs[1], (x - x.bfloat16().float()).bfloat16())
    y = torch.empty(M, N, device="cuda")
    engine.tc_linear_x3(xs[0], xs[1], ws[0], ws[1], b, None, y32=y)
    want = x.double() @ w.double().T + b.double()
    err = rel_err(y, want)
    assert err < 1e-4, err
    y1 = torch.empty(M, N, device="cuda")
    engine.tc_linear(xs[0], ws[0], b, None, y32=y1)
    assert rel_err(y1, want) > err          # the compensation is doing something


def test_cast_and_transpose(cuda_device):
    from pytorch_r2d2_b200 import engine
    x = torch.randn(1000, 1568, device="cuda")
    o = torch.empty(1000, 1568, device="cuda", dtype=torch.bfloat16)
    engine.cast_bf16(x, o)
    assert torch.equal(o, x.bfloat16())
    x2 = torch.randn(77, 130, device="cuda")
    o2 = torch.empty(130, 77, device="cuda", dtype=torch.bfloat16)
    engine.transpose_cast_bf16(x2, o2)
    assert torch.equal(o2, x2.T.contiguous().bfloat16())
