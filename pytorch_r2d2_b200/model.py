"""``QNet`` behind the reference's interface (/root/reference model.py:6-83): same constructor
argument, ``forward(state, return_hs_cs=False)``, ``reset`` / ``set_state`` / ``get_state``, same
``state_dict`` keys and default initialisation, module-held hidden state that carries over
between calls.  The conv torso, LSTM unroll and dueling head run as hand-written CUDA kernels
through the C ABI; parameters are ``nn.Parameter`` views into one flat f32 buffer.

Differences a caller can see: `device` must be a CUDA device (no CPU path); the width is a
constructor argument (``cell_size``, default 256 as in model.py:24); hs/cs are copied to host
numpy only when ``return_hs_cs=True`` (the reference always copies, model.py:68-69, and then
drops them).  Autograd works through ``forward`` (backward = our CUDA kernels); the incoming
hidden state is treated as a constant, which is how the reference uses it (learner.py:75-86).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib, engine
from .engine import Weights
from .params import FEAT, MODULE_ORDER, N_ACTIONS, ParamLayout


class _Group(nn.Module):
    """Holder giving parameters the reference's dotted names (e.g. ``vis_layers.0.weight``)."""


def _reference_order_init(cell: int):
    """Default init drawn in the construction order of model.py:12-36 (same RNG stream as
    ``QNet()`` in the reference for a given ``torch.manual_seed``)."""
    convs = [nn.Conv2d(4, 32, 8, 4), nn.Conv2d(32, 32, 4, 2), nn.Conv2d(32, 32, 3, 1)]
    lstm = nn.LSTMCell(FEAT, cell)
    val0, val2 = nn.Linear(cell, cell), nn.Linear(cell, 1)
    adv0, adv2 = nn.Linear(cell, cell), nn.Linear(cell, N_ACTIONS)
    out = {}
    for i, c in zip((0, 2, 4), convs):
        out[f"vis_layers.{i}.weight"], out[f"vis_layers.{i}.bias"] = c.weight.data, c.bias.data
    for k in ("weight_ih", "weight_hh", "bias_ih", "bias_hh"):
        out[f"lstm.{k}"] = getattr(lstm, k).data
    for name, l0, l2 in (("val", val0, val2), ("adv", adv0, adv2)):
        out[f"{name}.0.weight"], out[f"{name}.0.bias"] = l0.weight.data, l0.bias.data
        out[f"{name}.2.weight"], out[f"{name}.2.bias"] = l2.weight.data, l2.bias.data
    return out


class NetStash:
    """Activations of one forward pass (what backward needs)."""
    __slots__ = ("frames_u8", "frames_f32", "s0", "T", "B", "a1", "a2", "feat", "xproj", "hs_ext",
                 "cs_ext", "gates", "hid", "q")


def alloc_forward_ws(w: Weights, T, B, device, *, want_gates=True, head_from=0, tc=False):
    """Every buffer one ``net_forward`` call writes (activations, LSTM histories, head, and for
    the tensor-core path the bf16 operand images and the recurrence scratch).  Callers that run
    the same shape repeatedly (actors) allocate it once and pass it as ``ws``; nothing is then
    allocated per call, which also makes the call capturable in a CUDA graph."""
    H, R = w.H, T * B
    dev = device
    f = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)  # noqa: E731
    h16 = lambda *shape: torch.empty(shape, dtype=torch.bfloat16, device=dev)  # noqa: E731
    Mh = (T - head_from) * B
    ws = dict(key=(T, B, H, want_gates, head_from, tc))
    ws["a1"], ws["a2"], ws["feat"] = engine.alloc_torso(R, dev)
    ws["hs_ext"], ws["cs_ext"] = f(T + 1, B, H), f(T + 1, B, H)
    ws["gates"] = f(T, B, 4 * H) if want_gates else None
    ws["hid"], ws["q"] = f(Mh, 2 * H), f(Mh, N_ACTIONS)
    if tc:
        Rp = (R + 3) // 4 * 4                      # tc_linear needs N % 4 == 0 (N = frames here)
        ws["xproj"] = f(4 * H, Rp)                                   # transposed projection
        ws["w16"], ws["hw16"] = h16(2, 4 * H, FEAT), h16(2, 2 * H, H)
        ws["f16"] = torch.zeros((2, Rp, FEAT), dtype=torch.bfloat16, device=dev)   # pad rows stay zero
        ws["hd16"] = h16(2, Mh, H)
        ws["scratch"] = torch.empty(engine.tc_lstm_scratch_bytes(B, H), dtype=torch.uint8, device=dev)
    else:
        ws["xproj"] = f(R, 4 * H)
    return ws


def net_forward(w: Weights, T, B, h0, c0, *, slab=None, s0=0, frames=None, want_gates=True,
                head_from=0, ws=None, tc=False):
    """Torso -> input projection -> LSTM unroll -> head on steps [head_from, T).

    Frames come either from a u8 slab [B,S,4,84,84] (steps s0..s0+T, /255 fused) or from an f32
    tensor [T*B,4,84,84].  tc=True runs every contraction on tcgen05 tensor cores (compensated
    bf16x3, fp32-accurate forward).  ``ws`` = a workspace from ``alloc_forward_ws`` for this shape
    (default: allocated here).  Returns a NetStash whose tensors live in the workspace; q covers
    rows head_from*B .. T*B."""
    dev, H = h0.device, w.H
    R = T * B
    key = (T, B, H, want_gates, head_from, tc)
    if ws is None:
        ws = alloc_forward_ws(w, T, B, dev, want_gates=want_gates, head_from=head_from, tc=tc)
    elif ws["key"] != key:
        raise ValueError(f"workspace was allocated for {ws['key']}, this call is {key}")
    st = NetStash()
    st.frames_u8, st.frames_f32, st.s0, st.T, st.B = slab, frames, s0, T, B
    st.a1, st.a2, st.feat = ws["a1"], ws["a2"], ws["feat"]
    if slab is not None:
        engine.torso_fwd_u8(slab, s0, T, w, st.a1, st.a2, st.feat, tc=tc)
    else:
        engine.torso_fwd_f32(frames, w, st.a1, st.a2, st.feat, tc=tc)
    st.hs_ext, st.cs_ext = ws["hs_ext"], ws["cs_ext"]
    if h0.data_ptr() != st.hs_ext[0].data_ptr():
        st.hs_ext[0].copy_(h0)
    if c0.data_ptr() != st.cs_ext[0].data_ptr():
        st.cs_ext[0].copy_(c0)
    st.gates = ws["gates"]
    Mh = (T - head_from) * B
    st.hid, st.q = ws["hid"], ws["q"]
    hs_head = st.hs_ext[1 + head_from:].reshape(Mh, H)
    st.xproj = ws["xproj"]
    if tc:
        w16, f16, hd16, hw16 = ws["w16"], ws["f16"], ws["hd16"], ws["hw16"]
        engine.split_bf16(w.w_ih, w16[0], w16[1])
        engine.split_bf16(st.feat, f16[0, :R], f16[1, :R])
        engine.tc_linear_x3(w16[0], w16[1], f16[0], f16[1], w.b_ih, w.b_hh, y32=st.xproj, bias_row=True)
        engine.tc_lstm_seq_fwd(st.xproj, 0, w.w_hh, st.hs_ext, st.cs_ext, st.gates, T, B, H, ws["scratch"])
        engine.split_bf16(hs_head, hd16[0], hd16[1])
        engine.split_bf16(w.head_w1, hw16[0], hw16[1])
        engine.tc_linear_x3(hd16[0], hd16[1], hw16[0], hw16[1], w.head_b1, None, y32=st.hid, relu=True)
        engine.head_out(st.hid, w, st.q)
    else:
        engine.linear_fwd(st.feat, w.w_ih, w.b_ih, w.b_hh, st.xproj)
        engine.lstm_seq_fwd(st.xproj, w.w_hh, st.hs_ext, st.cs_ext, st.gates, T, B, H)
        engine.head_fwd(hs_head, w, st.hid, st.q)
    return st


def net_backward(w: Weights, g: Weights, st: NetStash, dq, t0=0):
    """Backward of ``net_forward`` for steps [t0, T) given dL/dq for those steps ([ (T-t0)*B, 6]).
    BPTT is truncated at t0 (no gradient into the incoming state).  Writes every gradient
    tensor of `g` (a Weights view over a flat gradient buffer); nothing is accumulated."""
    dev, H, B = dq.device, w.H, st.B
    L = st.T - t0
    M = L * B
    assert st.hid.shape[0] == M and dq.shape == (M, N_ACTIONS)
    dhid = torch.empty_like(st.hid)
    engine.head_bwd(dq, st.hid, w, dhid, g.head_w2, g.head_b2)
    hs_learn = st.hs_ext[1 + t0:].reshape(M, H)
    nbytes = max(engine.linear_wgrad_scratch_bytes(M, 2 * H, H),
                 engine.linear_wgrad_scratch_bytes(M, 4 * H, H),
                 engine.linear_wgrad_scratch_bytes(M, 4 * H, FEAT),
                 engine.torso_bwd_scratch_bytes(M))
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    engine.linear_wgrad(dhid, hs_learn, g.head_w1, g.head_b1, scratch)
    dhs = torch.empty((M, H), device=dev)
    engine.linear_dgrad(dhid, w.head_w1, dhs)
    dgates = torch.empty((M, 4 * H), device=dev)
    dc_work = torch.empty((B, H), device=dev)
    engine.lstm_seq_bwd(dhs, w.w_hh, st.gates[t0:], st.cs_ext[t0:], dgates, dc_work, L, B, H)
    hprev = st.hs_ext[t0:t0 + L].reshape(M, H)
    engine.linear_wgrad(dgates, hprev, g.w_hh, g.b_hh, scratch)
    g.b_ih.copy_(g.b_hh)
    feat_l = st.feat[t0 * B:]
    engine.linear_wgrad(dgates, feat_l, g.w_ih, None, scratch)
    dfeat = torch.empty((M, FEAT), device=dev)
    engine.linear_dgrad(dgates, w.w_ih, dfeat)
    a1, a2 = st.a1[t0 * B:], st.a2[t0 * B:]
    if st.frames_u8 is not None:
        engine.torso_bwd_u8(st.frames_u8, st.s0 + t0, L, a1, a2, feat_l, dfeat, w, g, scratch)
    else:
        engine.torso_bwd_f32(st.frames_f32[t0 * B:], a1, a2, feat_l, dfeat, w, g, scratch)


class _QNetFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, net, frames, T, B, h0, c0, *params):
        w = net._weights()
        st = net_forward(w, T, B, h0, c0, frames=frames, want_gates=True)
        ctx.net, ctx.st = net, st
        ctx.mark_non_differentiable(st.hs_ext, st.cs_ext)
        return st.q, st.hs_ext, st.cs_ext

    @staticmethod
    def backward(ctx, dq, _dh, _dc):
        net, st = ctx.net, ctx.st
        w = net._weights()
        gflat = torch.empty_like(net._flat)
        g = Weights(gflat, net.layout)
        net_backward(w, g, st, dq.contiguous(), 0)
        grads = tuple(net.layout.view(gflat, k) for k in MODULE_ORDER)
        return (None, None, None, None, None, None) + grads


class QNet(nn.Module):
    def __init__(self, device="cuda:0", cell_size=256, precision="fp32"):
        super().__init__()
        if precision not in ("fp32", "bf16"):
            raise ValueError("precision must be 'fp32' or 'bf16'")
        # "bf16": no-grad forwards (actors, target evaluation) run on tensor cores with the
        # compensated bf16x3 contractions; forwards that need autograd use the fp32 kernels
        self.precision = precision
        if not torch.cuda.is_available():
            raise RuntimeError("pytorch_r2d2_b200.QNet needs a CUDA device (no CPU fallback)")
        _lib.load()
        dev = torch.device(device)
        if dev.type != "cuda":
            raise ValueError(f"QNet runs on CUDA devices only, got {device!r}")
        self.device = dev
        self.cell_size = cell_size
        self.hs, self.cs = None, None
        self.layout = ParamLayout(cell_size)
        init = _reference_order_init(cell_size)
        flat = torch.empty(self.layout.total, dtype=torch.float32, device=dev)
        for k in MODULE_ORDER:
            self.layout.view(flat, k).copy_(init[k])
        self._flat = flat
        self._w = None
        groups = {}
        for top in ("vis_layers", "val", "adv"):
            groups[top] = nn.ModuleDict()
        self.vis_layers, self.lstm = groups["vis_layers"], _Group()
        self.val, self.adv = groups["val"], groups["adv"]
        for k in MODULE_ORDER:
            parts = k.split(".")
            p = nn.Parameter(self.layout.view(flat, k))
            if parts[0] == "lstm":
                self.lstm.register_parameter(parts[1], p)
            else:
                md = groups[parts[0]]
                if parts[1] not in md:
                    md[parts[1]] = _Group()
                md[parts[1]].register_parameter(parts[2], p)

    # ------------------------------------------------------------------ flat buffer upkeep
    def _named(self):
        d = dict(self.named_parameters())
        return [d[k] for k in MODULE_ORDER]

    def _flat_ok(self):
        base = self._flat.data_ptr()
        return all(p.data_ptr() == base + 4 * self.layout.offset(k) and p.device == self._flat.device
                   for k, p in zip(MODULE_ORDER, self._named()))

    def _reflatten(self, flat=None):
        """Re-point the parameters into one flat buffer (after .to()/deepcopy broke the views), or
        into a caller-supplied one (`flat`: e.g. the IPC-mapped buffer of the peer exchange)."""
        ps = self._named()
        dev = ps[0].device
        if dev.type != "cuda":
            raise RuntimeError("QNet parameters were moved off the GPU; there is no CPU path")
        if flat is None:
            flat = torch.empty(self.layout.total, dtype=torch.float32, device=dev)
        elif flat.numel() != self.layout.total or flat.dtype != torch.float32 or flat.device != dev:
            raise ValueError("rebase: need a contiguous f32 buffer of layout.total elements on the net's device")
        for k, p in zip(MODULE_ORDER, ps):
            self.layout.view(flat, k).copy_(p.data)
            p.data = self.layout.view(flat, k)
        self._flat, self._w, self.device = flat, None, dev

    def rebase(self, flat: torch.Tensor) -> None:
        """Move the parameters (values kept) into the given flat f32 buffer."""
        self._reflatten(flat)

    def _weights(self) -> Weights:
        if not self._flat_ok():
            self._reflatten()
        if self._w is None:
            self._w = Weights(self._flat, self.layout)
        return self._w

    def flat_parameters(self) -> torch.Tensor:
        self._weights()
        return self._flat

    def __deepcopy__(self, memo):
        new = QNet.__new__(QNet)
        nn.Module.__init__(new)
        src_flat = self.flat_parameters()
        new.device, new.cell_size, new.layout = self.device, self.cell_size, self.layout
        new.precision = self.precision
        new.hs, new.cs, new._w = None, None, None
        new._flat = src_flat.clone()
        groups = {top: nn.ModuleDict() for top in ("vis_layers", "val", "adv")}
        new.vis_layers, new.lstm = groups["vis_layers"], _Group()
        new.val, new.adv = groups["val"], groups["adv"]
        for k in MODULE_ORDER:
            parts = k.split(".")
            p = nn.Parameter(new.layout.view(new._flat, k))
            if parts[0] == "lstm":
                new.lstm.register_parameter(parts[1], p)
            else:
                md = groups[parts[0]]
                if parts[1] not in md:
                    md[parts[1]] = _Group()
                md[parts[1]].register_parameter(parts[2], p)
        return new

    def cpu_state_dict(self):
        """CPU copy of ``state_dict()``: what the reference publishes to its actors and saves
        (``deepcopy(net).cpu().state_dict()``, learner.py:119,123-124) without cloning the
        network on the device first.  The reference's own idiom keeps working as it is:
        ``deepcopy(net)`` is a detached device copy and ``nn.Module.cpu()`` on that copy moves
        its parameters to the host (such a copy can be read, not run: there is no CPU path)."""
        return {k: v.detach().cpu() for k, v in self.state_dict().items()}

    # ------------------------------------------------------------------ reference API
    def reset(self):
        self.hs, self.cs = None, None

    def set_state(self, hs, cs):
        self.hs, self.cs = hs, cs

    def get_state(self):
        return self.hs.detach().cpu().numpy(), self.cs.detach().cpu().numpy()

    def forward(self, state, return_hs_cs=False):
        if state.dim() == 5:
            T, B = int(state.shape[0]), int(state.shape[1])
        else:
            T, B = 1, int(state.shape[0])
        dev = self._flat.device
        frames = state.to(device=dev, dtype=torch.float32).reshape(T * B, 4, 84, 84).contiguous()
        H = self.cell_size
        if self.hs is None:
            self.hs = torch.zeros(B, H, device=dev)
            self.cs = torch.zeros(B, H, device=dev)
        h0 = torch.as_tensor(self.hs).to(device=dev, dtype=torch.float32).detach().contiguous()
        c0 = torch.as_tensor(self.cs).to(device=dev, dtype=torch.float32).detach().contiguous()
        w = self._weights()
        need_grad = torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters())
        if need_grad:
            q, hs_ext, cs_ext = _QNetFunction.apply(self, frames, T, B, h0, c0, *self._named())
        else:
            st = net_forward(w, T, B, h0, c0, frames=frames, want_gates=False,
                             tc=self.precision == "bf16")
            q, hs_ext, cs_ext = st.q, st.hs_ext, st.cs_ext
        self.hs, self.cs = hs_ext[T], cs_ext[T]
        q_val = q.view(-1, N_ACTIONS) if T > 1 else q
        if return_hs_cs:
            hs_seq = hs_ext[1:].reshape(T * B, H).detach().cpu().numpy()
            cs_seq = cs_ext[1:].reshape(T * B, H).detach().cpu().numpy()
            return q_val, hs_seq, cs_seq
        return q_val
