"""GPU parity: network kernels, QNet and the full Learner.train() step (through the C ABI)
against the oracle and the golden vectors of the unmodified reference.
Tolerance (fp32 mode): max|a-b| / max|b| <= 1e-5 per tensor (BASELINE.json north_star)."""
import os

import numpy as np
import pytest
import torch

from oracle import net_oracle as NO
from util import gpu_mem_np, make_oracle_replay, mirror_to_gpu, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-5
# stored priorities after 100 optimiser steps, max |gpu - oracle| / (|oracle| + 1e-3)
# (measured: fp32 0.077, bf16 0.29 -- (|delta|+eps)^0.6 of small TD errors is ill-conditioned)
PRIO_DRIFT = {"fp32": 0.1, "bf16": 0.4}
# median / max over a tensor of |w_gpu - w_oracle| / max|w| after 100 optimiser steps, and the
# Q-values of the two final networks on a probe batch.  Measured: fp32 1.8e-3 / 0.077 / 9.5e-3,
# bf16 5.2e-3 / 0.13 / 6.2e-2: per-step gradients are within 2e-2 (4e-3 measured) but centered
# RMSprop divides by sqrt(v - m^2), which is ~0.22 |g| whatever the size of g, so every weight
# moves ~lr/0.22 per step in the direction of a gradient SIGN that rounding can flip.
MEDIAN_DRIFT = {"fp32": 2.5e-3, "bf16": 8e-3}
MAX_DRIFT = {"fp32": 0.1, "bf16": 0.3}      # bf16: 0.13 / 0.20 measured with two summation orders
Q_DRIFT = {"fp32": 2e-2, "bf16": 1e-1}


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "network.npz"))


def make_net(H, seed=0):
    from pytorch_r2d2_b200 import QNet
    torch.manual_seed(seed)
    return QNet("cuda:0", H)


def test_init_and_state_dict_match_reference_layout(cuda_device):
    net = make_net(256)
    p = NO.init_params(0, 256)
    sd = net.state_dict()
    assert list(sd.keys()) == list(NO.PARAM_KEYS)
    for k in NO.PARAM_KEYS:
        assert torch.equal(sd[k].cpu(), p[k]), k
    assert sum(x.numel() for x in net.parameters()) == 2037095


def test_torso_linear_vs_torch(cuda_device):
    from pytorch_r2d2_b200 import engine
    net = make_net(256)
    w = net._weights()
    p = {k: v.cpu() for k, v in net.state_dict().items()}
    rng = np.random.RandomState(1)
    B, S = 3, 5
    slab = torch.from_numpy(rng.randint(0, 256, (B, S, 4, 84, 84)).astype(np.uint8)).cuda()
    a1, a2, feat = engine.alloc_torso(B * 3, "cuda")
    engine.torso_fwd_u8(slab, 1, 3, w, a1, a2, feat)          # frames 1..3, time-major rows
    frames = slab[:, 1:4].permute(1, 0, 2, 3, 4).reshape(-1, 4, 84, 84).cpu().float() / 255.
    want = NO.torso(p, frames)
    assert rel_err(feat, want) < TOL
    x1 = torch.relu(torch.nn.functional.conv2d(frames, p["vis_layers.0.weight"], p["vis_layers.0.bias"], stride=4))
    assert rel_err(a1, x1) < TOL
    # f32 entry point gives the same numbers
    a1b, a2b, featb = engine.alloc_torso(B * 3, "cuda")
    engine.torso_fwd_f32(frames.cuda().contiguous(), w, a1b, a2b, featb)
    assert rel_err(featb, want) < TOL
    # input projection with both biases
    xp = torch.empty((B * 3, 4 * 256), device="cuda")
    engine.linear_fwd(feat, w.w_ih, w.b_ih, w.b_hh, xp)
    wantxp = want @ p["lstm.weight_ih"].T + p["lstm.bias_ih"] + p["lstm.bias_hh"]
    assert rel_err(xp, wantxp) < TOL


@pytest.mark.parametrize("M,N,K", [(7, 5, 3), (64, 1024, 256), (640, 2048, 1568), (130, 70, 33)])
def test_linear_ops_vs_torch(cuda_device, M, N, K):
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(M * N + K)
    x, w, dy = torch.randn(M, K, generator=g), torch.randn(N, K, generator=g), torch.randn(M, N, generator=g)
    b = torch.randn(N, generator=g)
    xd, wd, dyd, bd = x.cuda(), w.cuda(), dy.cuda(), b.cuda()
    y = torch.empty(M, N, device="cuda")
    engine.linear_fwd(xd, wd, bd, None, y, relu=True)
    assert rel_err(y, torch.relu(x.double() @ w.double().T + b.double())) < TOL
    dx = torch.empty(M, K, device="cuda")
    engine.linear_dgrad(dyd, wd, dx)
    assert rel_err(dx, dy.double() @ w.double()) < TOL
    engine.linear_dgrad(dyd, wd, dx, accumulate=True)
    assert rel_err(dx, 2 * (dy.double() @ w.double())) < TOL
    nb = engine.linear_wgrad_scratch_bytes(M, N, K)
    scratch = torch.empty(max(nb, 16), dtype=torch.uint8, device="cuda")
    dw, db = torch.empty(N, K, device="cuda"), torch.empty(N, device="cuda")
    engine.linear_wgrad(dyd, xd, dw, db, scratch)
    assert rel_err(dw, dy.double().T @ x.double()) < TOL
    assert rel_err(db, dy.double().sum(0)) < TOL


@pytest.mark.parametrize("H,B,T", [(256, 4, 5), (512, 64, 3), (256, 70, 2)])
def test_lstm_fwd_bwd_vs_autograd(cuda_device, H, B, T):
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(H + B)
    xproj = torch.randn(T, B, 4 * H, generator=g) * 0.5
    w_hh = torch.randn(4 * H, H, generator=g) * 0.05
    h0, c0 = torch.randn(B, H, generator=g) * 0.1, torch.randn(B, H, generator=g) * 0.1
    dhs = torch.randn(T, B, H, generator=g)
    # torch reference
    xr, wr = xproj.clone().requires_grad_(True), w_hh.clone().requires_grad_(True)
    h, c, hs, cs = h0, c0, [], []
    for t in range(T):
        gates = xr[t] + h @ wr.T
        i, f, gg, o = gates.chunk(4, 1)
        c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
        h = torch.sigmoid(o) * torch.tanh(c)
        hs.append(h)
        cs.append(c)
    hs_t = torch.stack(hs)
    (hs_t * dhs).sum().backward()
    # ours
    hs_ext = torch.empty(T + 1, B, H, device="cuda")
    cs_ext = torch.empty(T + 1, B, H, device="cuda")
    hs_ext[0], cs_ext[0] = h0.cuda(), c0.cuda()
    gates_d = torch.empty(T, B, 4 * H, device="cuda")
    engine.lstm_seq_fwd(xproj.cuda(), w_hh.cuda(), hs_ext, cs_ext, gates_d, T, B, H)
    assert rel_err(hs_ext[1:], hs_t.detach()) < TOL
    assert rel_err(cs_ext[1:], torch.stack(cs).detach()) < TOL
    dgates = torch.empty(T, B, 4 * H, device="cuda")
    dc = torch.empty(B, H, device="cuda")
    engine.lstm_seq_bwd(dhs.cuda(), w_hh.cuda(), gates_d, cs_ext, dgates, dc, T, B, H)
    assert rel_err(dgates, xr.grad) < TOL                      # dL/dxproj == pre-activation grads
    hprev = hs_ext[:T].reshape(T * B, H)
    dw = torch.empty(4 * H, H, device="cuda")
    nb = engine.linear_wgrad_scratch_bytes(T * B, 4 * H, H)
    engine.linear_wgrad(dgates.reshape(T * B, 4 * H), hprev, dw, None,
                        torch.empty(max(nb, 16), dtype=torch.uint8, device="cuda"))
    assert rel_err(dw, wr.grad) < TOL


@pytest.mark.parametrize("H,M", [(256, 16), (512, 300)])
def test_head_fwd_bwd_vs_autograd(cuda_device, H, M):
    from pytorch_r2d2_b200 import engine
    net = make_net(H, seed=3)
    w = net._weights()
    p = {k: v.cpu().clone().requires_grad_(True) for k, v in net.state_dict().items()}
    g = torch.Generator().manual_seed(5)
    hs = torch.randn(M, H, generator=g).requires_grad_(True)
    dq = torch.randn(M, 6, generator=g)
    q = NO.head(p, hs)
    (q * dq).sum().backward()
    hid = torch.empty(M, 2 * H, device="cuda")
    qd = torch.empty(M, 6, device="cuda")
    engine.head_fwd(hs.detach().cuda(), w, hid, qd)
    assert rel_err(qd, q.detach()) < TOL
    gflat = torch.zeros_like(net.flat_parameters())
    gw = engine.Weights(gflat, net.layout)
    dhid = torch.empty_like(hid)
    engine.head_bwd(dq.cuda(), hid, w, dhid, gw.head_w2, gw.head_b2)
    assert rel_err(gw.head_w2[0], p["val.2.weight"].grad[0]) < TOL
    assert rel_err(gw.head_w2[1:], p["adv.2.weight"].grad) < TOL
    assert rel_err(gw.head_b2[1:], p["adv.2.bias"].grad) < TOL
    assert rel_err(gw.head_b2[:1], p["val.2.bias"].grad) < TOL
    dhs = torch.empty(M, H, device="cuda")
    engine.linear_dgrad(dhid, w.head_w1, dhs)
    assert rel_err(dhs, hs.grad) < TOL


def test_torso_bwd_vs_autograd(cuda_device):
    from pytorch_r2d2_b200 import engine
    net = make_net(256, seed=4)
    w = net._weights()
    p = {k: v.cpu().clone().requires_grad_(True) for k, v in net.state_dict().items()}
    rng = np.random.RandomState(8)
    B, S, s0, nt = 3, 6, 2, 3
    slab = torch.from_numpy(rng.randint(0, 256, (B, S, 4, 84, 84)).astype(np.uint8)).cuda()
    R = B * nt
    frames = slab[:, s0:s0 + nt].permute(1, 0, 2, 3, 4).reshape(R, 4, 84, 84).cpu().float() / 255.
    feat_ref = NO.torso(p, frames)
    dfeat = torch.randn(R, 1568, generator=torch.Generator().manual_seed(1))
    (feat_ref * dfeat).sum().backward()
    a1, a2, feat = engine.alloc_torso(R, "cuda")
    engine.torso_fwd_u8(slab, s0, nt, w, a1, a2, feat)
    gflat = torch.zeros_like(net.flat_parameters())
    gw = engine.Weights(gflat, net.layout)
    scratch = torch.empty(engine.torso_bwd_scratch_bytes(R), dtype=torch.uint8, device="cuda")
    engine.torso_bwd_u8(slab, s0, nt, a1, a2, feat, dfeat.cuda(), w, gw, scratch)
    for name, t in (("vis_layers.4.weight", gw.w3), ("vis_layers.4.bias", gw.b3),
                    ("vis_layers.2.weight", gw.w2), ("vis_layers.2.bias", gw.b2),
                    ("vis_layers.0.weight", gw.w1), ("vis_layers.0.bias", gw.b1)):
        assert rel_err(t, p[name].grad) < TOL, name


def test_qnet_forward_vs_golden(cuda_device, gold):
    """QNet.forward == the reference's forward: 5-D zero state, 4-D set_state, carry-over."""
    net = make_net(256)
    rng = np.random.RandomState(int(gold["fwd/seed_x"]))
    x5 = torch.from_numpy(rng.randint(0, 256, (3, 2, 4, 84, 84)).astype(np.float32) / np.float32(255)).cuda()
    x4 = x5[0].clone()
    h0 = torch.from_numpy((rng.randn(2, 256) * 0.1).astype(np.float32)).cuda()
    c0 = torch.from_numpy((rng.randn(2, 256) * 0.1).astype(np.float32)).cuda()
    with torch.no_grad():
        net.reset()
        q5, hs5, cs5 = net(x5, True)
        net.set_state(h0, c0)
        q4, hs4, cs4 = net(x4, True)
        q4b, hs4b, cs4b = net(x4, True)
    assert q5.shape == (6, 6) and q4.shape == (2, 6) and isinstance(hs5, np.ndarray)
    for got, key in ((q5, "q5"), (hs5, "hs5"), (cs5, "cs5"), (q4, "q4"), (hs4, "hs4"), (cs4, "cs4"),
                     (q4b, "q4b"), (hs4b, "hs4b"), (cs4b, "cs4b")):
        assert rel_err(got, gold["fwd/" + key]) < TOL, key
    h, c = net.get_state()
    assert rel_err(h, gold["fwd/hs4b"]) < TOL


def test_qnet_autograd_vs_oracle(cuda_device):
    """loss.backward() through QNet.forward populates .grad like the reference module."""
    net = make_net(256, seed=5)
    p = {k: v.cpu().clone() for k, v in net.state_dict().items()}
    rng = np.random.RandomState(11)
    x = torch.from_numpy(rng.randint(0, 256, (4, 3, 4, 84, 84)).astype(np.float32) / np.float32(255))
    h0 = torch.from_numpy((rng.randn(3, 256) * 0.1).astype(np.float32))
    c0 = torch.from_numpy((rng.randn(3, 256) * 0.1).astype(np.float32))
    coef = torch.randn(12, 6, generator=torch.Generator().manual_seed(2))
    pr = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    q_ref, _, _, _ = NO.qnet_forward(pr, x, h0, c0)
    (q_ref * coef).sum().backward()
    net.set_state(h0.cuda(), c0.cuda())
    q = net(x.cuda())
    (q * coef.cuda()).sum().backward()
    assert rel_err(q, q_ref.detach()) < TOL
    got = dict(net.named_parameters())
    for k in NO.PARAM_KEYS:
        assert rel_err(got[k].grad, pr[k].grad) < TOL, k


def _learner(H, M, B, BI, LN, NS, seed=0, precision="fp32"):
    from pytorch_r2d2_b200 import Learner
    torch.manual_seed(seed)
    L = Learner(1, {}, device="cuda:0", cell_size=H, memory_size=M, precision=precision)
    L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = B, BI, LN, NS
    L.sequence_length = BI + LN
    return L


def test_train_step_vs_golden_reference(cuda_device, gold):
    """Three full Learner.train() steps against the unmodified reference's outputs."""
    M, B, BI, LN, NS, H = [int(v) for v in gold["train/config"]]
    rep, _ = make_oracle_replay(M, H, NS, BI, LN, seed=21)
    L = _learner(H, M, B, BI, LN, NS)
    L.replay_memory = mirror_to_gpu(rep, B)
    for step in range(3):
        torch.manual_seed(1000 + step)
        L.train()
        if step == 0:
            out = L.last
            assert np.array_equal(out["seq_index"].cpu().numpy(), gold["train/seq_index"])   # bit-exact
            assert rel_err(out["q_all"], gold["train/q_all"]) < TOL
            assert rel_err(out["q_next_online"], gold["train/q_next_online"]) < TOL
            assert rel_err(out["q_next_target"], gold["train/q_next_target"]) < TOL
            gflat = L.optim.grad
            for k in NO.PARAM_KEYS:
                f = L.net.layout.view(gflat, k).reshape(-1)
                stp = max(1, f.numel() // 2048)
                err = float((f[::stp][:2048].cpu() - torch.from_numpy(gold[f"train/grad_sub/{k}"])).abs().max()
                            / float(gold[f"train/grad_absmax/{k}"]))
                assert err < TOL, (k, err)
                assert abs(float(f.double().sum()) - float(gold[f"train/grad_sum/{k}"])) \
                    <= 1e-4 * max(1.0, float(gold[f"train/grad_absmax/{k}"]) * f.numel() ** 0.5), k
            sd = L.net.state_dict()
            for k in NO.PARAM_KEYS:
                f = sd[k].reshape(-1)
                stp = max(1, f.numel() // 2048)
                assert rel_err(f[::stp][:2048], gold[f"train/param1_sub/{k}"]) < TOL, k
                sq = L.net.layout.view(L.optim.square_avg, k).reshape(-1)
                assert rel_err(sq[::stp][:2048], gold[f"train/sqavg_sub/{k}"]) < 1e-4, k
            assert np.allclose(gpu_mem_np(L.replay_memory, "priority"), gold["train/priority_after"],
                               rtol=1e-4, atol=1e-7)
            assert np.allclose(gpu_mem_np(L.replay_memory, "sequence_priority"),
                               gold["train/seqprio_after"], rtol=1e-4, atol=1e-7)
    sd = L.net.state_dict()
    for k in NO.PARAM_KEYS:                                     # drift after 3 optimiser steps
        f = sd[k].reshape(-1)
        stp = max(1, f.numel() // 2048)
        assert rel_err(f[::stp][:2048], gold[f"train/param3_sub/{k}"]) < 1e-4, k
    assert np.allclose(gpu_mem_np(L.replay_memory, "priority"), gold["train/priority_after3"], rtol=1e-3, atol=1e-6)


@pytest.mark.parametrize("H,B,BI,LN,NS", [(512, 8, 4, 4, 5), (256, 5, 2, 3, 1)])
def test_train_step_vs_oracle(cuda_device, H, B, BI, LN, NS):
    """Wider net / other lengths: one step against the oracle run on the same batch, targets
    compared given the oracle's a* (argmax ties are discontinuous), flip rate reported."""
    M = 80
    rep, _ = make_oracle_replay(M, H, NS, BI, LN, seed=40 + H)
    L = _learner(H, M, B, BI, LN, NS, seed=7)
    L.replay_memory = mirror_to_gpu(rep, B)
    p = {k: v.cpu().clone() for k, v in L.net.state_dict().items()}
    pt = {k: v.cpu().clone() for k, v in L.target_net.state_dict().items()}
    # ReLU' is discontinuous: a conv pre-activation within rounding distance of 0 (seen: 4e-8)
    # legitimately flips one mask bit between two fp32 implementations and changes that
    # channel's gradient by O(1e-3).  Draw the batch so that no learn-window pre-activation of
    # the torso sits closer than 1e-6 to the kink (deterministic search over seeds).
    for attempt in range(20):
        u = np.random.RandomState(H + 1000 * attempt).rand(B)
        si = rep.sample_indices(u)
        tb = NO.to_torch_batch(rep.gather(si)["batch"])
        if NO.torso_kink_margin(p, tb["state"][BI:].reshape(-1, 4, 84, 84)) > 1e-6:
            break
    res = NO.train_step(p, pt, tb, burn_in=BI, n_step=NS)
    L.train(u=torch.from_numpy(u), next_action_override=res["next_action"].cuda())
    out = L.last
    assert np.array_equal(out["seq_index"].cpu().numpy(), si)
    flips = int((out["next_action"].cpu() != res["next_action"]).sum())
    assert flips == 0
    for k in ("q_all", "q_next_online", "q_next_target", "y", "delta"):
        assert rel_err(out[k], res[k]) < TOL, k
    assert rel_err(out["loss"], res["loss"]) < TOL
    assert rel_err(out["priority"], res["priority"]) < 1e-4
    for k in NO.PARAM_KEYS:
        assert rel_err(L.net.layout.view(L.optim.grad, k), res["grads"][k]) < TOL, k


def test_value_rescale_flag(cuda_device):
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(3)
    n = 50
    q, qo, qt = (torch.randn(n, 6, generator=g) * 3 for _ in range(3))
    a = torch.randint(0, 6, (n,), generator=g)
    r, d = torch.randn(n, generator=g), (torch.rand(n, generator=g) < 0.2).float()
    na = qo.argmax(1)
    y = NO.value_rescale(r + 0.99 ** 5 * NO.value_rescale_inv(qt.gather(1, na[:, None])[:, 0]) * (1 - d))
    out = dict(next_action=torch.empty(n, dtype=torch.int64, device="cuda"), y=torch.empty(n, device="cuda"),
               delta=torch.empty(n, device="cuda"), dq=torch.empty(n, 6, device="cuda"),
               prio=torch.empty(n, device="cuda"), loss=torch.empty(1, device="cuda"))
    engine.td_target(q.cuda(), qo.cuda(), qt.cuda(), a.cuda(), r.cuda(), d.cuda(), 0.99 ** 5, out, rescale=True)
    assert torch.equal(out["next_action"].cpu(), na)
    assert rel_err(out["y"], y) < 1e-5


def test_rmsprop_vs_torch(cuda_device):
    from pytorch_r2d2_b200 import engine
    g = torch.Generator().manual_seed(0)
    n = 1003
    p0, grads = torch.randn(n, generator=g), [torch.randn(n, generator=g) * 0.01 for _ in range(4)]
    pr = p0.clone().requires_grad_(True)
    opt = torch.optim.RMSprop([pr], lr=0.00025 / 4.0, alpha=0.95, eps=1.5e-7, centered=True)
    pd = p0.cuda()
    sq, ga = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    for gr in grads:
        pr.grad = gr.clone()
        opt.step()
        engine.rmsprop_centered(pd, gr.cuda(), sq, ga, 0.00025 / 4.0, 0.95, 1.5e-7)
    assert rel_err(pd, pr.detach()) < 1e-6
    assert rel_err(sq, opt.state[pr]["square_avg"]) < 1e-5


def test_full_size_step_properties(cuda_device):
    """BASELINE shape (B=64, T=40+40, n=5, H=512): the step runs, the loss is finite, gradients
    are finite and non-zero, and two identical steps from identical state give identical
    results (determinism; no atomics on the data path)."""
    M, B, BI, LN, NS, H = 4000, 64, 40, 40, 5, 512
    outs = []
    for rep_i in range(2):
        L = _learner(H, M, B, BI, LN, NS, seed=1)
        rm = L.replay_memory
        gen = torch.Generator(device="cuda").manual_seed(5)
        raw = lambda k: dict.__getitem__(rm.memory, k)  # noqa: E731
        raw("state").random_(0, 256, generator=gen)
        raw("hs_cs").normal_(0, 0.1, generator=gen)
        raw("target_hs_cs").normal_(0, 0.1, generator=gen)
        raw("action").random_(0, 6, generator=gen)
        raw("reward").normal_(0, 1, generator=gen)
        raw("priority").uniform_(0.01, 1.01, generator=gen)
        rm.index = M
        rm.burn_in_length, rm.learning_length, rm.sequence_length = BI, LN, BI + LN
        starts = torch.arange(0, M - 85, 40, device="cuda")
        rm.update_sequence_priority(starts)
        rm.memory["is_seq_start"][starts] = 1
        u = torch.rand(B, dtype=torch.float64, generator=torch.Generator().manual_seed(9))
        loss = L.train(u=u)
        torch.cuda.synchronize()
        g = L.optim.grad
        assert torch.isfinite(loss).all() and torch.isfinite(g).all() and float(g.abs().max()) > 0
        outs.append((loss.clone(), g.clone(), L.net.flat_parameters().clone(),
                     raw("sequence_priority").clone()))
    for a, b in zip(outs[0], outs[1]):
        assert torch.equal(a, b)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_cuda_graph_step_equals_eager(cuda_device, precision):
    """Replaying the captured step gives the same weights / priorities as enqueueing it."""
    M, B, BI, LN, NS, H = 96, 8, 3, 4, 2, 256
    outs = []
    for graphed in (False, True):
        rep, _ = make_oracle_replay(M, H, NS, BI, LN, seed=5)
        from pytorch_r2d2_b200 import Learner
        torch.manual_seed(3)
        L = Learner(1, {}, device="cuda:0", cell_size=H, memory_size=M, precision=precision)
        L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = B, BI, LN, NS
        L.replay_memory = mirror_to_gpu(rep, B)
        L.use_cuda_graph = graphed
        torch.manual_seed(77)
        losses = [float(L.train()) for _ in range(4)]
        torch.cuda.synchronize()
        outs.append((losses, L.net.flat_parameters().clone(), gpu_mem_np(L.replay_memory, "sequence_priority")))
    assert outs[0][0] == outs[1][0]
    assert torch.equal(outs[0][1], outs[1][1])
    assert np.array_equal(outs[0][2], outs[1][2])


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-5), ("bf16", 1e-5)])
def test_batched_actor_inference_vs_oracle(cuda_device, precision, tol):
    """BASELINE config 5 shape in miniature: N actors stepped together (T=1), both nets, hidden
    state carried per actor across steps == the reference's per-actor batch-1 loop
    (actor.py:96-106).  The bf16 mode (tensor cores, compensated) stays inside 1e-5 too."""
    from pytorch_r2d2_b200 import BatchedActor
    N, H = 37, 256
    torch.manual_seed(11)
    act = BatchedActor(N, "cuda:0", H, precision=precision)
    p = {k: v.cpu().clone() for k, v in act.net.state_dict().items()}
    rng = np.random.RandomState(4)
    h = c = th = tc = None
    for step in range(3):
        x = torch.from_numpy(rng.randint(0, 256, (N, 4, 84, 84)).astype(np.float32) / np.float32(255))
        with torch.no_grad():
            q_ref, _, _, (h, c) = NO.qnet_forward(p, x, h, c)
            tq_ref, _, _, (th, tc) = NO.qnet_forward(p, x, th, tc)
        action, q, _, _, tq, _, _ = act.select_action(x.cuda(), rng=np.random.RandomState(step))
        assert rel_err(q, q_ref) < tol and rel_err(tq, tq_ref) < tol, (step, rel_err(q, q_ref))
        assert action.shape == (N,) and int(action.min()) >= 0 and int(action.max()) < 6
    hh, cc = act.net.get_state()
    assert rel_err(hh, h) < tol and rel_err(cc, c) < tol
    prio = act.calc_priority(q.cpu().numpy(), action.cpu().numpy(), rng.randn(N).astype(np.float32),
                             q.cpu().numpy(), tq.cpu().numpy(), rng.rand(N) < 0.2)
    assert prio.shape == (N,) and np.all(prio > 0)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_hundred_step_drift_vs_oracle(cuda_device, precision):
    """100 consecutive learner steps (sample -> train -> RMSprop -> priority update)
    against the oracle doing the same 100 steps on the CPU.  Both sides draw from the same
    uniforms; the GPU is fed the oracle's indices (one ulp of priority difference must not fork
    the two trajectories) and its own tree must agree with them on the way."""
    M, B, BI, LN, NS, H = 64, (4 if precision == "fp32" else 8), 2, 3, 2, 256   # bf16: L*B % 8 == 0
    rep, _ = make_oracle_replay(M, H, NS, BI, LN, seed=33)
    L = _learner(H, M, B, BI, LN, NS, seed=3, precision=precision)
    L.replay_memory = mirror_to_gpu(rep, B)
    p = {k: v.cpu().clone() for k, v in L.net.state_dict().items()}
    pt = {k: v.cpu().clone() for k, v in L.target_net.state_dict().items()}
    sq = {k: torch.zeros_like(v) for k, v in p.items()}
    ga = {k: torch.zeros_like(v) for k, v in p.items()}
    rng = np.random.RandomState(5)
    agree = 0
    for step in range(100):
        u = rng.rand(B)
        si = rep.sample_indices(u)
        res = NO.train_step(p, pt, NO.to_torch_batch(rep.gather(si)["batch"]), burn_in=BI, n_step=NS)
        for k in NO.PARAM_KEYS:
            NO.rmsprop_centered(p[k], res["grads"][k], sq[k], ga[k])
        T = BI + LN
        idx = (si[None, :] + np.arange(BI, T)[:, None]) % M          # index[BI:] time-major
        rep.update_priority(idx.reshape(-1), np.asarray(res["priority"]).reshape(-1))
        rep.update_sequence_priority(si, True)
        own = L.replay_memory.sample_indices(torch.from_numpy(u)).cpu().numpy()
        agree += int(np.array_equal(own, si))
        L.train(seq_index=torch.from_numpy(si).cuda(), next_action_override=res["next_action"].cuda())
    # Centered RMSprop moves every weight by ~lr/0.22 per step whatever the size of its gradient
    # (v - m^2 = a(1-a) g^2 after one step), so weights whose gradient is at rounding-noise level
    # legitimately random-walk apart between two fp32 implementations.  What must agree: the bulk
    # of every tensor, and the function the network computes.
    sd = L.net.state_dict()
    stats = {}
    for k in NO.PARAM_KEYS:
        d = (sd[k].cpu() - p[k]).abs().reshape(-1) / float(p[k].abs().max())
        stats[k] = (float(d.median()), float(d.quantile(0.99)) if d.numel() > 1 else float(d.max()), float(d.max()))
    probe = NO.to_torch_batch(rep.gather(rep.sample_indices(np.random.RandomState(9).rand(B)))["batch"])
    want = NO.train_step(p, pt, probe, burn_in=BI, n_step=NS)
    mine = NO.train_step({k: v.cpu() for k, v in sd.items()}, pt, probe, burn_in=BI, n_step=NS)
    qerr = rel_err(mine["q_all"], want["q_all"])
    gp, op = gpu_mem_np(L.replay_memory, "priority"), rep.memory["priority"]
    pdev = float(np.max(np.abs(gp - op) / (np.abs(op) + 1e-3)))
    print(f"[{precision}] drift after 100 steps: q {qerr:.2e}, priority {pdev:.2e}, tree agreement {agree}/100, "
          f"worst tensor median {max(v[0] for v in stats.values()):.2e} max {max(v[2] for v in stats.values()):.2e}")
    for k in NO.PARAM_KEYS:
        assert stats[k][0] < MEDIAN_DRIFT[precision], (k, stats[k])
        assert stats[k][2] < MAX_DRIFT[precision], (k, stats[k])
    assert qerr < Q_DRIFT[precision], (qerr, stats)
    assert pdev < PRIO_DRIFT[precision], (pdev, stats)
    # the GPU tree draws the same sequences as the oracle's on (almost) every step; in bf16 mode its
    # priorities drift further from the oracle's (measured 96 and 92 of 100 with two equally valid
    # summation orders of the conv2 kernel), so a draw lands on the other side of a CDF boundary
    # a little more often
    assert agree >= (95 if precision == "fp32" else 85), agree
