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
    # masked form: outputs where mask <= 0 are exactly zero, the others equal the plain product
    mask = torch.randn(M, N, generator=g).cuda()
    y32p, y32m = torch.empty(M, N, device="cuda"), torch.full((M, N), float("nan"), device="cuda")
    engine.tc_linear(x, w, None, None, y32=y32p)
    engine.tc_linear_masked(x, w, mask, y32m)
    assert torch.equal(y32m, torch.where(mask > 0, y32p, torch.zeros_like(y32p)))


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


def test_tc_linear_x3_wide_forms_agree(cuda_device):
    """The input-projection shape (wide tiles): the CTA-pair kernel with multicast B tiles, the
    same with a CTA cap (max_ctas: the kernel is persistent, fewer CTAs walk more tiles each) and an
    odd M-tile count (single-CTA kernel) give the fp32-accurate product; pair and capped runs are
    bit-identical (same per-tile arithmetic)."""
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(77)
    for M, N, K in ((2048, 5440, 1568), (1920, 5000, 200)):      # 16 and 15 M-tiles
        x = torch.randn(M, K, generator=g).cuda()
        w = (torch.randn(N, K, generator=g) * 0.1).cuda()
        b = torch.randn(M, generator=g).cuda()
        xs = torch.empty(2, M, K, device="cuda", dtype=torch.bfloat16)
        ws = torch.empty(2, N, K, device="cuda", dtype=torch.bfloat16)
        engine.split_bf16(x, xs[0], xs[1])
        engine.split_bf16(w, ws[0], ws[1])
        want = x.double() @ w.double().T + b.double()[:, None]
        ys = []
        for cap in (0, 84, 2):
            y = torch.full((M, N), float("nan"), device="cuda")
            engine.tc_linear_x3(xs[0], xs[1], ws[0], ws[1], b, None, y32=y, bias_row=True, max_ctas=cap)
            assert rel_err(y, want) < 1e-4, (M, cap, rel_err(y, want))
            ys.append(y)
        assert torch.equal(ys[0], ys[1]) and torch.equal(ys[0], ys[2])


@pytest.mark.parametrize("Kc,P,Q", [(2560, 2048, 512), (2560, 2048, 1568), (200, 72, 40), (64, 128, 128), (130, 8, 264)])
def test_tc_linear_tn_vs_fp64(cuda_device, Kc, P, Q):
    """y = a^T . b with both operands read MN-major (no transposed copies) == the fp64 product of
    the same bf16 values to fp32-accumulation accuracy, incl. ragged K, P and Q."""
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(Kc + P + Q)
    a = torch.randn(Kc, P, generator=g).bfloat16().cuda()
    b = torch.randn(Kc, Q, generator=g).bfloat16().cuda()
    y = torch.full((P, Q), float("nan"), device="cuda")
    engine.tc_linear_tn(a, b, y)
    want = a.double().T @ b.double()
    assert rel_err(y, want) < 1e-5, rel_err(y, want)     # fp32 accumulation over Kc terms


@pytest.mark.parametrize("M,N,K", [(2560, 1568, 2048), (2560, 512, 1024), (100, 72, 40), (128, 128, 64), (130, 264, 8)])
def test_tc_linear_nn_vs_fp64(cuda_device, M, N, K):
    """y = a . b with b [K, N] as stored (read MN-major) == the fp64 product of the same bf16 values."""
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(M + N + K)
    a = torch.randn(M, K, generator=g).bfloat16().cuda()
    b = torch.randn(K, N, generator=g).bfloat16().cuda()
    y = torch.full((M, N), float("nan"), device="cuda")
    engine.tc_linear_nn(a, b, y)
    want = a.double() @ b.double()
    assert rel_err(y, want) < 1e-5, rel_err(y, want)


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


def test_tc_torso_vs_torch(cuda_device):
    """Tensor-core conv torso (bf16x3 implicit GEMM) == fp32 torch convs to 1e-5, u8 slab and
    f32 entry points, incl. a partial last tile and the time-major frame mapping."""
    import numpy as np
    from oracle import net_oracle as NO
    from pytorch_r2d2_b200 import QNet, engine
    torch.manual_seed(2)
    net = QNet("cuda:0", 256)
    w = net._weights()
    p = {k: v.cpu() for k, v in net.state_dict().items()}
    rng = np.random.RandomState(1)
    B, S = 3, 5
    slab = torch.from_numpy(rng.randint(0, 256, (B, S, 4, 84, 84)).astype(np.uint8)).cuda()
    a1, a2, feat = engine.alloc_torso(B * 3, "cuda")
    engine.torso_fwd_u8(slab, 1, 3, w, a1, a2, feat, tc=True)
    frames = slab[:, 1:4].permute(1, 0, 2, 3, 4).reshape(-1, 4, 84, 84).cpu().float() / 255.
    x1 = torch.relu(torch.nn.functional.conv2d(frames, p["vis_layers.0.weight"], p["vis_layers.0.bias"], stride=4))
    x2 = torch.relu(torch.nn.functional.conv2d(x1, p["vis_layers.2.weight"], p["vis_layers.2.bias"], stride=2))
    want = NO.torso(p, frames)
    assert rel_err(a1, x1) < 1e-5, rel_err(a1, x1)
    assert rel_err(a2, x2) < 1e-5, rel_err(a2, x2)
    assert rel_err(feat, want) < 1e-5, rel_err(feat, want)
    b1, b2, featb = engine.alloc_torso(B * 3, "cuda")
    engine.torso_fwd_f32(frames.cuda().contiguous(), w, b1, b2, featb, tc=True)
    assert rel_err(featb, want) < 1e-5
    # many tiles per CTA (exercises the ring / TMEM double buffering across > 148*2 tiles)
    # (960 frames: every conv layer has several tiles per persistent CTA, which is what exposes
    #  ring-phase hazards between the four MMA-issuing warps)
    slab2 = torch.from_numpy(rng.randint(0, 256, (32, 30, 4, 84, 84)).astype(np.uint8)).cuda()
    c1, c2, f2 = engine.alloc_torso(32 * 30, "cuda")
    d1, d2, g2 = engine.alloc_torso(32 * 30, "cuda")
    for _ in range(3):
        engine.torso_fwd_u8(slab2, 0, 30, w, c1, c2, f2, tc=True)
    engine.torso_fwd_u8(slab2, 0, 30, w, d1, d2, g2, tc=False)
    assert rel_err(c1, d1) < 1e-5 and rel_err(c2, d2) < 1e-5 and rel_err(f2, g2) < 1e-5


def test_tc_conv1_dual_matches_single(cuda_device):
    """One pass of the first conv layer for two nets (online: frames 0..S, target: frames s0..S)
    == the two single-net passes (gather form; the dual pass is the shifted-window kernel, which
    sums in another order: equal to fp32 rounding, and to 1e-5 against an fp64 conv of the same
    frames), plus the tail entry point (conv2 + conv3)."""
    import numpy as np
    from pytorch_r2d2_b200 import QNet, engine
    torch.manual_seed(5)
    na, nb = QNet("cuda:0", 256), QNet("cuda:0", 256)
    wa, wb = na._weights(), nb._weights()
    rng = np.random.RandomState(4)
    for B, S, s0 in ((3, 6, 2), (32, 30, 5)):
        slab = torch.from_numpy(rng.randint(0, 256, (B, S, 4, 84, 84)).astype(np.uint8)).cuda()
        a1, a2, fa = engine.alloc_torso(B * S, "cuda")
        b1, b2, fb = engine.alloc_torso(B * (S - s0), "cuda")
        engine.torso_fwd_u8(slab, 0, S, wa, a1, a2, fa, tc=True)
        engine.torso_fwd_u8(slab, s0, S - s0, wb, b1, b2, fb, tc=True)
        c1, c2, fc = engine.alloc_torso(B * S, "cuda")
        d1, d2, fd = engine.alloc_torso(B * (S - s0), "cuda")
        d1.fill_(float("nan"))
        engine.conv1_fwd_dual_u8(slab, S, wa, c1, wb, d1, s0)
        assert rel_err(c1, a1) < 5e-6 and rel_err(d1, b1) < 5e-6, (rel_err(c1, a1), rel_err(d1, b1))
        assert not torch.isnan(d1).any()
        fr = slab.transpose(0, 1).reshape(B * S, 4, 84, 84).double() / 255.0     # rows are time-major
        c1_ref = torch.relu(torch.nn.functional.conv2d(fr, wa.w1.double(), wa.b1.double(), stride=4))
        d1_ref = torch.relu(torch.nn.functional.conv2d(slab[:, s0:].transpose(0, 1).reshape(-1, 4, 84, 84).double() / 255.0,
                                                       wb.w1.double(), wb.b1.double(), stride=4))
        assert rel_err(c1, c1_ref) < 1e-5 and rel_err(d1, d1_ref) < 1e-5, (rel_err(c1, c1_ref), rel_err(d1, d1_ref))
        f16 = torch.zeros(2, B * S, 1568, dtype=torch.bfloat16, device="cuda")
        engine.torso_fwd_tail(c1, wa, c2, fc, f16)
        engine.torso_fwd_tail(d1, wb, d2, fd)
        # the tail runs the shifted-window kernels (other summation order than the gather form):
        # same values to fp32 rounding, and to 1e-5 against torch's convs
        # (fp64: torch's fp32 convs on the GPU default to TF32)
        a2_ref = torch.relu(torch.nn.functional.conv2d(c1.double(), wa.w2.double(), wa.b2.double(), stride=2))
        f_ref = torch.relu(torch.nn.functional.conv2d(a2_ref, wa.w3.double(), wa.b3.double(), stride=1)).reshape(B * S, 1568)
        assert rel_err(c2, a2_ref) < 1e-5 and rel_err(fc, f_ref) < 1e-5, (rel_err(c2, a2_ref), rel_err(fc, f_ref))
        assert rel_err(c2, a2) < 5e-6 and rel_err(fc, fa) < 5e-6 and rel_err(fd, fb) < 5e-6
        # the fused (hi, lo) image of feat == r2d2_split_bf16(feat)
        hi = fc.bfloat16()
        assert torch.equal(f16[0], hi) and torch.equal(f16[1], (fc - hi.float()).bfloat16())


@pytest.mark.parametrize("B,nt", [(1, 1), (3, 1), (2, 5)])
def test_tc_conv_backward_few_frames(cuda_device, B, nt):
    """Shifted-window conv gradients with fewer frames than CTAs (persistent CTAs without work
    write zero partials) and frame counts that are not a multiple of anything: == the fp32 kernels
    to the bf16 class."""
    import numpy as np
    from pytorch_r2d2_b200 import QNet, engine
    torch.manual_seed(9)
    net = QNet("cuda:0", 256)
    w = net._weights()
    S, s0 = nt + 1, 1
    slab = torch.from_numpy(np.random.RandomState(B * 10 + nt).randint(0, 256, (B, S, 4, 84, 84)).astype(np.uint8)).cuda()
    R = B * nt
    a1, a2, feat = engine.alloc_torso(R, "cuda")
    engine.torso_fwd_u8(slab, s0, nt, w, a1, a2, feat)
    dfeat = torch.randn(R, 1568, generator=torch.Generator().manual_seed(2)).cuda()
    outs = []
    for fn in (engine.torso_bwd_u8, engine.torso_bwd_tc_u8):
        gflat = torch.full_like(net.flat_parameters(), float("nan"))
        gw = engine.Weights(gflat, net.layout)
        scratch = torch.empty(engine.torso_bwd_scratch_bytes(R), dtype=torch.uint8, device="cuda")
        fn(slab, s0, nt, a1, a2, feat, dfeat.clone(), w, gw, scratch)
        outs.append(gw)
    ref, tc = outs
    for name in ("w1", "b1", "w2", "b2", "w3", "b3"):
        err = rel_err(getattr(tc, name), getattr(ref, name))
        assert err < 1e-2, (name, err)


@pytest.mark.parametrize("H,B,T", [(256, 4, 5), (512, 64, 6), (512, 8, 3), (256, 100, 2)])
def test_tc_lstm_vs_torch(cuda_device, H, B, T):
    """Persistent tensor-core LSTM unroll == fp32 recurrence to 1e-5 (bf16x3 compensation),
    incl. batches that are not a multiple of the 64-column tile and > 1 batch group."""
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(H + B + T)
    xproj = torch.randn(T, B, 4 * H, generator=g) * 0.5
    w_hh = torch.randn(4 * H, H, generator=g) * 0.05
    h0, c0 = torch.randn(B, H, generator=g) * 0.1, torch.randn(B, H, generator=g) * 0.1
    h, c, hs, cs, gs = h0, c0, [], [], []
    for t in range(T):
        gates = xproj[t] + h @ w_hh.T
        i, f, gg, o = gates.chunk(4, 1)
        i, f, gg, o = torch.sigmoid(i), torch.sigmoid(f), torch.tanh(gg), torch.sigmoid(o)
        c = f * c + i * gg
        h = o * torch.tanh(c)
        hs.append(h); cs.append(c); gs.append(torch.cat([i, f, gg, o], 1))
    xprojT = xproj.reshape(T * B, 4 * H).T.contiguous().cuda()          # [4H, T*B]
    pad = torch.zeros(4 * H, 3 * B, device="cuda")                        # exercise col0 != 0
    xprojT = torch.cat([pad, xprojT], 1).contiguous()
    hs_ext = torch.empty(T + 1, B, H, device="cuda"); cs_ext = torch.empty(T + 1, B, H, device="cuda")
    hs_ext[0], cs_ext[0] = h0.cuda(), c0.cuda()
    gates_d = torch.empty(T, B, 4 * H, device="cuda")
    scratch = torch.empty(engine.tc_lstm_scratch_bytes(B, H), dtype=torch.uint8, device="cuda")
    t0 = 1                                                                # split image from step 1 on
    h16 = torch.zeros(2, T - t0, B, H, dtype=torch.bfloat16, device="cuda")
    for _ in range(2):                                                    # twice: scratch reuse
        engine.tc_lstm_seq_fwd(xprojT, 3 * B, w_hh.cuda(), hs_ext, cs_ext, gates_d, T, B, H, scratch,
                               h16=(h16[0], h16[1]), hist_t0=t0)
    torch.cuda.synchronize()
    hi = hs_ext[1 + t0:].bfloat16()
    assert torch.equal(h16[0], hi) and torch.equal(h16[1], (hs_ext[1 + t0:] - hi.float()).bfloat16())
    assert rel_err(hs_ext[1:], torch.stack(hs)) < 1e-5, rel_err(hs_ext[1:], torch.stack(hs))
    assert rel_err(cs_ext[1:], torch.stack(cs)) < 1e-5
    assert rel_err(gates_d, torch.stack(gs)) < 1e-5


@pytest.mark.parametrize("H,B,T", [(256, 4, 5), (512, 64, 6), (256, 100, 3)])
def test_tc_lstm_bwd_vs_autograd(cuda_device, H, B, T):
    """Persistent tensor-core BPTT == autograd of the fp32 recurrence within the bf16 class
    (plain bf16 operands in the backward contraction): 1e-2 relative to max|grad|."""
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(H * 3 + B + T)
    xproj = (torch.randn(T, B, 4 * H, generator=g) * 0.5).requires_grad_(True)
    w_hh = torch.randn(4 * H, H, generator=g) * 0.05
    h0, c0 = torch.randn(B, H, generator=g) * 0.1, torch.randn(B, H, generator=g) * 0.1
    dhs = torch.randn(T, B, H, generator=g)
    h, c, hs, cs, gs = h0, c0, [], [c0], []
    for t in range(T):
        gates = xproj[t] + h @ w_hh.T
        i, f, gg, o = gates.chunk(4, 1)
        i, f, gg, o = torch.sigmoid(i), torch.sigmoid(f), torch.tanh(gg), torch.sigmoid(o)
        c = f * c + i * gg
        h = o * torch.tanh(c)
        hs.append(h); cs.append(c); gs.append(torch.cat([i, f, gg, o], 1))
    (torch.stack(hs) * dhs).sum().backward()
    gates_d = torch.stack(gs).detach().cuda().contiguous()
    cs_ext = torch.stack(cs).detach().cuda().contiguous()
    dgates = torch.full((T, B, 4 * H), float("nan"), device="cuda")
    scratch = torch.empty(engine.tc_lstm_bwd_scratch_bytes(B, H), dtype=torch.uint8, device="cuda")
    dg16 = torch.zeros(T, B, 4 * H, dtype=torch.bfloat16, device="cuda")
    for _ in range(2):
        engine.tc_lstm_seq_bwd(dhs.cuda(), w_hh.cuda(), gates_d, cs_ext, dgates, T, B, H, scratch, dgates16=dg16)
    torch.cuda.synchronize()
    err = rel_err(dgates, xproj.grad)
    assert err < 1e-2, err
    assert torch.equal(dg16, dgates.bfloat16())              # the fused bf16 copy
    assert rel_err(dgates[-1], xproj.grad[-1]) < 1e-5        # last step has no recurrent term


def test_tc_conv_backward_vs_fp32(cuda_device):
    """Tensor-core conv backward (data + weight gradients in plain bf16) vs torch autograd of the
    fp32 convs AND vs the fp32 kernels on the same inputs: 1e-2 relative to max|grad| (bf16
    operand rounding)."""
    import numpy as np
    from oracle import net_oracle as NO
    from pytorch_r2d2_b200 import QNet, engine
    torch.manual_seed(4)
    net = QNet("cuda:0", 256)
    w = net._weights()
    rng = np.random.RandomState(8)
    B, S, s0, nt = 5, 7, 2, 4
    slab = torch.from_numpy(rng.randint(0, 256, (B, S, 4, 84, 84)).astype(np.uint8)).cuda()
    R = B * nt
    a1, a2, feat = engine.alloc_torso(R, "cuda")
    engine.torso_fwd_u8(slab, s0, nt, w, a1, a2, feat)
    dfeat = torch.randn(R, 1568, generator=torch.Generator().manual_seed(1)).cuda()
    outs = []
    for fn in (engine.torso_bwd_u8, engine.torso_bwd_tc_u8):
        gflat = torch.full_like(net.flat_parameters(), float("nan"))
        gw = engine.Weights(gflat, net.layout)
        scratch = torch.empty(engine.torso_bwd_scratch_bytes(R), dtype=torch.uint8, device="cuda")
        fn(slab, s0, nt, a1, a2, feat, dfeat.clone(), w, gw, scratch)
        outs.append(gw)
    ref, tc = outs
    # directly against autograd (the oracle's torso on the same frames)
    p = {k: v.cpu().clone().requires_grad_(True) for k, v in net.state_dict().items()}
    frames = slab[:, s0:s0 + nt].permute(1, 0, 2, 3, 4).reshape(R, 4, 84, 84).cpu().float() / 255.
    (NO.torso(p, frames) * dfeat.cpu()).sum().backward()
    for name, key in (("w1", "vis_layers.0.weight"), ("b1", "vis_layers.0.bias"), ("w2", "vis_layers.2.weight"),
                      ("b2", "vis_layers.2.bias"), ("w3", "vis_layers.4.weight"), ("b3", "vis_layers.4.bias")):
        err = rel_err(getattr(tc, name), p[key].grad)
        assert err < 1e-2, (name, err)
        assert rel_err(getattr(ref, name), p[key].grad) < 1e-5, name
    for name in ("w1", "w2", "w3"):
        err = rel_err(getattr(tc, name), getattr(ref, name))
        assert err < 1e-2, (name, err)
    for name in ("b1", "b2", "b3"):   # data gradients now also run in bf16 on tensor cores
        assert rel_err(getattr(tc, name), getattr(ref, name)) < 1e-2, name


def test_conv1_ring_forms_match_slab_forms(cuda_device):
    """First conv layer reading the sampled frames in place from the replay ring (incl. sequences
    that wrap around the end of the ring) == the same kernels on the gathered slab, bit for bit."""
    import numpy as np
    from pytorch_r2d2_b200 import QNet, _lib, engine
    torch.manual_seed(6)
    na, nb = QNet("cuda:0", 256), QNet("cuda:0", 256)
    wa, wb = na._weights(), nb._weights()
    rng = np.random.RandomState(8)
    Mring, B, S, s0 = 50, 5, 9, 2
    ring = torch.from_numpy(rng.randint(0, 256, (Mring, 4, 84, 84)).astype(np.uint8)).cuda()
    seq = torch.tensor([0, 13, 44, 49, 13], dtype=torch.int64, device="cuda")     # two wrap, one duplicate
    rows = (seq[:, None] + torch.arange(S, device="cuda")[None, :]) % Mring
    slab = ring[rows].contiguous()                                                # [B,S,4,84,84]
    a1, _, _ = engine.alloc_torso(B * S, "cuda")
    b1, _, _ = engine.alloc_torso(B * (S - s0), "cuda")
    c1, _, _ = engine.alloc_torso(B * S, "cuda")
    d1, _, _ = engine.alloc_torso(B * (S - s0), "cuda")
    engine.conv1_fwd_dual_u8(slab, S, wa, a1, wb, b1, s0)
    bits = torch.zeros(B * S, 400, dtype=torch.int32, device="cuda")
    engine.conv1_fwd_dual_ring(ring, seq, S, wa, c1, wb, d1, s0, relu_bits_a=bits)
    assert torch.equal(a1, c1) and torch.equal(b1, d1)
    # ReLU bit mask of net a: bit co of word (frame, pixel) == (a1[frame, co, pixel] > 0)
    want_bits = ((c1.reshape(B * S, 32, 400) > 0).long() << torch.arange(32, device="cuda").view(1, 32, 1)).sum(1)
    assert torch.equal(bits.long() & 0xFFFFFFFF, want_bits)
    # conv2 data gradient with ReLU' from the bit mask == from the activations
    R = B * S
    dz2 = torch.randn(R, 32, 9, 9, device="cuda")
    g_act, g_bits = torch.empty(R, 32, 20, 20, device="cuda"), torch.full((R, 32, 20, 20), float("nan"), device="cuda")
    _lib.call("r2d2_tc_conv_dgrad2", _lib.ptr(dz2, "f32"), _lib.ptr(wa.w2), _lib.ptr(c1, "f32"), _lib.ptr(g_act, "f32"),
              R, _lib.stream_ptr())
    _lib.call("r2d2_tc_conv_dgrad2_masked", _lib.ptr(dz2, "f32"), _lib.ptr(wa.w2), _lib.ptr(bits, "i32"),
              _lib.ptr(g_bits, "f32"), R, _lib.stream_ptr())
    # (the bit-mask form is the shifted-window kernel: same products, another summation order)
    assert rel_err(g_bits, g_act) < 1e-6 and torch.equal(g_act == 0, g_bits == 0)
    want = torch.nn.grad.conv2d_input((R, 32, 20, 20), wa.w2.double(), dz2.double(), stride=2) * (c1 > 0)
    assert rel_err(g_bits, want) < 1e-2          # bf16 operands
    # weight gradient of the first layer over the window [s0, S)
    nt = S - s0
    dz = torch.randn(B * nt, 32, 20, 20, device="cuda")
    part = torch.empty(_lib.load().r2d2_sm_count() * (32 * 256 + 32), device="cuda")
    g_slab, g_ring = torch.empty(32, 4, 8, 8, device="cuda"), torch.empty(32, 4, 8, 8, device="cuda")
    b_slab, b_ring = torch.empty(32, device="cuda"), torch.empty(32, device="cuda")
    _lib.call("r2d2_tc_conv_wgrad_layer", 1, _lib.ptr(slab, "u8"), B, S, s0, nt, _lib.ptr(dz, "f32"),
              _lib.ptr(g_slab, "f32"), _lib.ptr(b_slab, "f32"), _lib.ptr(part, "f32"), _lib.stream_ptr())
    engine.conv1_wgrad_ring(ring, seq, s0, nt, dz, g_ring, part, db1=b_ring)
    assert torch.equal(g_slab, g_ring) and torch.equal(b_slab, b_ring)
    # the fused bias gradient == the per-channel sum of dz
    want = dz.double().sum(dim=(0, 2, 3))
    assert rel_err(b_ring, want) < 1e-5
