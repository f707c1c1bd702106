"""Thin tensor-level wrappers over the C ABI (one Python function per entry point) and the
forward/backward compositions shared by ``QNet`` and ``Learner``.  No math happens here:
every function validates tensors and enqueues CUDA kernels on torch's current stream.
"""
from __future__ import annotations

import torch

from . import _lib
from .params import FEAT, N_ACTIONS

P = _lib.ptr
S = _lib.stream_ptr


class Weights:
    """Pointers to the tensors of one network inside its flat buffer (see params.py)."""

    def __init__(self, flat: torch.Tensor, layout):
        v = lambda k: layout.view(flat, k)  # noqa: E731
        self.H = layout.cell
        self.w1, self.b1 = v("vis_layers.0.weight"), v("vis_layers.0.bias")
        self.w2, self.b2 = v("vis_layers.2.weight"), v("vis_layers.2.bias")
        self.w3, self.b3 = v("vis_layers.4.weight"), v("vis_layers.4.bias")
        self.w_ih, self.w_hh = v("lstm.weight_ih"), v("lstm.weight_hh")
        self.b_ih, self.b_hh = v("lstm.bias_ih"), v("lstm.bias_hh")
        H = self.H
        o, n = layout.span("val.0.weight", "adv.0.weight")
        self.head_w1 = flat[o:o + n].view(2 * H, H)
        o, n = layout.span("val.0.bias", "adv.0.bias")
        self.head_b1 = flat[o:o + n]
        o, n = layout.span("val.2.weight", "adv.2.weight")
        self.head_w2 = flat[o:o + n].view(1 + N_ACTIONS, H)
        o, n = layout.span("val.2.bias", "adv.2.bias")
        self.head_b2 = flat[o:o + n]


# ---------------------------------------------------------------------------- forward ops
def torso_fwd_u8(slab, s0, nt, w: Weights, a1, a2, feat, tc=False):
    B, Sfr = slab.shape[0], slab.shape[1]
    _lib.call("r2d2_tc_torso_fwd_u8" if tc else "r2d2_torso_fwd_u8", P(slab, "u8"), B, Sfr, s0, nt, P(w.w1), P(w.b1), P(w.w2), P(w.b2),
              P(w.w3), P(w.b3), P(a1, "f32"), P(a2, "f32"), P(feat, "f32"), S())


def conv1_fwd_dual_u8(slab, nt, wa: Weights, a1_a, wb: Weights, a1_b, s0_b):
    """First conv layer of two nets over the same frames: net a on frames 0..nt, net b on s0_b..nt."""
    B, Sfr = slab.shape[0], slab.shape[1]
    _lib.call("r2d2_tc_conv1_fwd_dual_u8", P(slab, "u8"), B, Sfr, nt, P(wa.w1), P(wa.b1), P(a1_a, "f32"),
              P(wb.w1), P(wb.b1), P(a1_b, "f32"), s0_b, S())


def conv1_fwd_dual_ring(ring, seq_index, nt, wa: Weights, a1_a, wb: Weights, a1_b, s0_b, relu_bits_a=None):
    """As conv1_fwd_dual_u8, the frames read in place from the replay ring [M,4,84,84] u8 through
    the sampled start slots seq_index (i64 [B], device)."""
    _lib.call("r2d2_tc_conv1_fwd_dual_ring", P(ring, "u8"), ring.shape[0], P(seq_index, "i64"), seq_index.numel(), nt,
              P(wa.w1), P(wa.b1), P(a1_a, "f32"), P(wb.w1), P(wb.b1), P(a1_b, "f32"), s0_b,
              P(relu_bits_a, "i32", True), S())


def conv1_wgrad_ring(ring, seq_index, s0, nt, dz1, dw1, part, db1=None):
    """First-layer weight gradient (and, with db1, bias gradient) over frames s0..s0+nt of the
    sampled sequences read in place from the ring; part: sm_count*(32*256+32) floats."""
    _lib.call("r2d2_tc_conv1_wgrad_ring", P(ring, "u8"), ring.shape[0], P(seq_index, "i64"), seq_index.numel(), s0, nt,
              P(dz1, "f32"), P(dw1), P(db1, None, True), P(part, "f32"), S())


def torso_fwd_tail(a1, w: Weights, a2, feat, feat16=None):
    """conv2 + conv3 of one net from its first-layer activations; feat16 = (hi, lo) bf16 buffers
    that also receive the split image of feat."""
    hi, lo = (P(feat16[0], "bf16"), P(feat16[1], "bf16")) if feat16 is not None else (None, None)
    _lib.call("r2d2_tc_torso_fwd_tail", P(a1, "f32"), a1.shape[0], P(w.w2), P(w.b2), P(w.w3), P(w.b3),
              P(a2, "f32"), P(feat, "f32"), hi, lo, S())


def torso_fwd_f32(frames, w: Weights, a1, a2, feat, tc=False):
    R = frames.shape[0]
    _lib.call("r2d2_tc_torso_fwd_f32" if tc else "r2d2_torso_fwd_f32", P(frames, "f32"), R, P(w.w1), P(w.b1), P(w.w2), P(w.b2), P(w.w3),
              P(w.b3), P(a1, "f32"), P(a2, "f32"), P(feat, "f32"), S())


def linear_fwd(x, w, b1, b2, y, relu=False):
    M, K = x.shape
    N = w.shape[0]
    assert w.shape[1] == K and y.shape == (M, N)
    _lib.call("r2d2_linear_fwd", P(x, "f32"), P(w, "f32"), P(b1, "f32", True), P(b2, "f32", True),
              P(y, "f32"), M, N, K, int(relu), S())


def lstm_seq_fwd(xproj, w_hh, hs_ext, cs_ext, gates, T, B, H):
    _lib.call("r2d2_lstm_seq_fwd", P(xproj, "f32"), P(w_hh, "f32"), P(hs_ext, "f32"), P(cs_ext, "f32"),
              P(gates, "f32", True), T, B, H, S())


def head_fwd(hs, w: Weights, hid, q):
    M, H = hs.shape
    _lib.call("r2d2_head_fwd", P(hs, "f32"), P(w.head_w1), P(w.head_b1), P(w.head_w2), P(w.head_b2),
              P(hid, "f32"), P(q, "f32"), M, H, S())


def head_out(hid, w: Weights, q):
    M, H2 = hid.shape
    _lib.call("r2d2_head_out", P(hid, "f32"), P(w.head_w2), P(w.head_b2), P(q, "f32"), M, H2 // 2, S())


def colsum(dy, db, db_copy=None):
    M, N = dy.shape
    if db_copy is not None:
        _lib.call("r2d2_colsum2", P(dy, "f32"), P(db, "f32"), P(db_copy, "f32"), M, N, S())
    else:
        _lib.call("r2d2_colsum", P(dy, "f32"), P(db, "f32"), M, N, S())


def td_target(q_all, qn_on, qn_tg, action, reward, done, gamma_n, out, *, rescale=False,
              rescale_eps=1e-3, prio_eps=1e-6, alpha=0.6, next_action_override=None):
    n = q_all.shape[0]
    _lib.call("r2d2_td_target", P(q_all, "f32"), P(qn_on, "f32"), P(qn_tg, "f32"), P(action, "i64"),
              P(reward, "f32"), P(done, "f32"), n, float(gamma_n), int(rescale), float(rescale_eps),
              float(prio_eps), float(alpha), P(next_action_override, "i64", True),
              P(out["next_action"], "i64"), P(out["y"], "f32"), P(out["delta"], "f32"),
              P(out["dq"], "f32"), P(out["prio"], "f32"), P(out["loss"], "f32"), S())


# ---------------------------------------------------------------------------- backward ops
def head_bwd(dq, hid, w: Weights, dhid, dw2, db2):
    M, H2 = hid.shape
    opt = lambda t: P(t, "f32") if t is not None else None  # noqa: E731
    _lib.call("r2d2_head_bwd", P(dq, "f32"), P(hid, "f32"), P(w.head_w2), opt(dhid), opt(dw2),
              opt(db2), M, H2 // 2, S())


def linear_dgrad(dy, w, dx, accumulate=False):
    M, N = dy.shape
    K = w.shape[1]
    _lib.call("r2d2_linear_dgrad", P(dy, "f32"), P(w, "f32"), P(dx, "f32"), M, N, K, int(accumulate), S())


def linear_wgrad_scratch_bytes(M, N, K):
    return int(_lib.load().r2d2_linear_wgrad_scratch_bytes(M, N, K))


def linear_wgrad(dy, x, dw, db, scratch):
    M, N = dy.shape
    K = x.shape[1]
    _lib.call("r2d2_linear_wgrad", P(dy, "f32"), P(x, "f32"), P(dw, "f32"), P(db, "f32", True), M, N, K,
              P(scratch, None, True), S())


def lstm_seq_bwd(dhs, w_hh, gates, cs_ext, dgates, dc_work, T, B, H):
    _lib.call("r2d2_lstm_seq_bwd", P(dhs, "f32"), P(w_hh, "f32"), P(gates, "f32"), P(cs_ext, "f32"),
              P(dgates, "f32"), P(dc_work, "f32"), T, B, H, S())


def torso_bwd_scratch_bytes(R):
    return int(_lib.load().r2d2_torso_bwd_scratch_bytes(R))


def torso_bwd_u8(slab, s0, nt, a1, a2, feat, dfeat, w: Weights, g: Weights, scratch):
    B, Sfr = slab.shape[0], slab.shape[1]
    _lib.call("r2d2_torso_bwd_u8", P(slab, "u8"), B, Sfr, s0, nt, P(a1, "f32"), P(a2, "f32"),
              P(feat, "f32"), P(dfeat, "f32"), P(w.w2), P(w.w3), P(g.w1), P(g.b1), P(g.w2), P(g.b2),
              P(g.w3), P(g.b3), P(scratch), S())


def torso_bwd_tc_u8(slab, s0, nt, a1, a2, feat, dfeat, w: Weights, g: Weights, scratch):
    """Tensor-core conv backward: ReLU mask, data gradients (tcgen05, ReLU' fused), bias sums,
    then the three weight gradients (tcgen05).  scratch layout [dz2 | dz1 | partials]."""
    B, Sfr = slab.shape[0], slab.shape[1]
    R = B * nt
    sc = scratch.view(torch.float32)
    dz2, dz1 = sc[:R * 2592], sc[R * 2592: R * (2592 + 12800)]
    part = sc[R * (2592 + 12800):]
    st = S()
    _lib.call("r2d2_relu_mask", P(dfeat, "f32"), P(feat, "f32"), dfeat.numel(), st)
    _lib.call("r2d2_tc_conv_dgrad3", P(dfeat, "f32"), P(w.w3), P(a2, "f32"), P(dz2, "f32"), R, st)
    _lib.call("r2d2_tc_conv_dgrad2", P(dz2, "f32"), P(w.w2), P(a1, "f32"), P(dz1, "f32"), R, st)
    _lib.call("r2d2_chansum", P(dfeat, "f32"), P(g.b3), R, 49, P(part, "f32"), st)
    _lib.call("r2d2_chansum", P(dz2, "f32"), P(g.b2), R, 81, P(part, "f32"), st)
    _lib.call("r2d2_chansum", P(dz1, "f32"), P(g.b1), R, 400, P(part, "f32"), st)
    _lib.call("r2d2_tc_torso_wgrad_u8", P(slab, "u8"), B, Sfr, s0, nt, P(a1, "f32"), P(a2, "f32"),
              P(dz1, "f32"), P(dz2, "f32"), P(dfeat, "f32"), P(g.w1), P(g.w2), P(g.w3), P(part, "f32"), st)


def torso_bwd_f32(frames, a1, a2, feat, dfeat, w: Weights, g: Weights, scratch):
    R = frames.shape[0]
    _lib.call("r2d2_torso_bwd_f32", P(frames, "f32"), R, P(a1, "f32"), P(a2, "f32"), P(feat, "f32"),
              P(dfeat, "f32"), P(w.w2), P(w.w3), P(g.w1), P(g.b1), P(g.w2), P(g.b2), P(g.w3), P(g.b3),
              P(scratch), S())


def rmsprop_centered(param, grad, sq, ga, lr, alpha, eps, grad_scale=1.0):
    _lib.call("r2d2_rmsprop_centered", P(param, "f32"), P(grad, "f32"), P(sq, "f32"), P(ga, "f32"),
              param.numel(), float(lr), float(alpha), float(eps), float(grad_scale), S())


# ---------------------------------------------------------------------------- bf16 tensor-core ops
def tc_linear(x16, w16, b1, b2, y32=None, y16=None, relu=False):
    M, K = x16.shape
    N = w16.shape[0]
    assert w16.shape[1] == K
    _lib.call("r2d2_tc_linear", P(x16, "bf16"), P(w16, "bf16"), P(b1, "f32", True), P(b2, "f32", True),
              P(y32, "f32", True), P(y16, "bf16", True), M, N, K, int(relu), S())


def tc_linear_masked(x16, w16, mask, y32):
    """y = (mask > 0) * (x . w^T): data gradient with the downstream ReLU' fused."""
    M, K = x16.shape
    N = w16.shape[0]
    _lib.call("r2d2_tc_linear_masked", P(x16, "bf16"), P(w16, "bf16"), P(mask, "f32"), P(y32, "f32"), M, N, K, S())


def tc_linear_nn(a16, b16, y32):
    """y [M, N] = a . b for row-major bf16 a [M, K], b [K, N] (data gradients; the weight as stored)."""
    M, K = a16.shape
    N = b16.shape[1]
    assert b16.shape[0] == K and y32.numel() == M * N
    _lib.call("r2d2_tc_linear_nn", P(a16, "bf16"), P(b16, "bf16"), P(y32, "f32"), M, N, K, S())


def tc_linear_tn(a16, b16, y32):
    """y [P, Q] = a^T . b for row-major bf16 a [K, P], b [K, Q] (weight gradients; no transposes)."""
    Kc, Pn = a16.shape
    Qn = b16.shape[1]
    assert b16.shape[0] == Kc and y32.numel() == Pn * Qn
    _lib.call("r2d2_tc_linear_tn", P(a16, "bf16"), P(b16, "bf16"), P(y32, "f32"), Kc, Pn, Qn, S())


def tc_linear_x3(x_hi, x_lo, w_hi, w_lo, b1, b2, y32=None, y16=None, relu=False, bias_row=False, max_ctas=0):
    """max_ctas > 0: run on at most that many SMs (the kernel is persistent; see the header)."""
    M, K = x_hi.shape
    N = w_hi.shape[0]
    _lib.call("r2d2_tc_linear_x3_ctas", P(x_hi, "bf16"), P(x_lo, "bf16"), P(w_hi, "bf16"), P(w_lo, "bf16"),
              P(b1, "f32", True), P(b2, "f32", True), P(y32, "f32", True), P(y16, "bf16", True),
              M, N, K, int(relu) | (2 if bias_row else 0), int(max_ctas), S())


def tc_lstm_bwd_scratch_bytes(B, H):
    return int(_lib.load().r2d2_tc_lstm_bwd_scratch_bytes(B, H))


def tc_lstm_seq_bwd(dhs, w_hh, gates, cs_ext, dgates, T, B, H, scratch, dgates16=None):
    _lib.call("r2d2_tc_lstm_seq_bwd", P(dhs, "f32"), P(w_hh, "f32"), P(gates, "f32"), P(cs_ext, "f32"),
              P(dgates, "f32"), P(dgates16, "bf16", True), T, B, H, P(scratch), S())


def tc_lstm_scratch_bytes(B, H):
    return int(_lib.load().r2d2_tc_lstm_scratch_bytes(B, H))


def tc_lstm_seq_fwd(xprojT, col0, w_hh, hs_ext, cs_ext, gates, T, B, H, scratch, h16=None, hist_t0=0):
    """xprojT: [4H, ldx] f32 (transposed input projection, biases included).  h16 = (hi, lo) bf16
    [T - hist_t0, B, H] buffers that receive the split image of h_t for t >= hist_t0."""
    hi, lo = (P(h16[0], "bf16"), P(h16[1], "bf16")) if h16 is not None else (None, None)
    _lib.call("r2d2_tc_lstm_seq_fwd", P(xprojT, "f32"), xprojT.shape[1], col0, P(w_hh, "f32"),
              P(hs_ext, "f32"), P(cs_ext, "f32"), P(gates, "f32", True), hi, lo, hist_t0, T, B, H,
              P(scratch), S())


def split_bf16(x, hi, lo):
    _lib.call("r2d2_split_bf16", P(x, "f32"), P(hi, "bf16"), P(lo, "bf16"), x.numel(), S())


def cast_bf16(x, out):
    _lib.call("r2d2_cast_bf16", P(x, "f32"), P(out, "bf16"), x.numel(), S())


def transpose_cast_bf16(x, out):
    R, C = x.shape
    assert out.shape == (C, R)
    _lib.call("r2d2_transpose_cast_bf16", P(x, "f32"), P(out, "bf16"), R, C, S())


# ---------------------------------------------------------------------------- compositions
def alloc_torso(R, device):
    return (torch.empty((R, 32, 20, 20), device=device), torch.empty((R, 32, 9, 9), device=device),
            torch.empty((R, FEAT), device=device))
