"""Oracle (CPU restatement, torch fp32) of the reference's network and train step.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.  Citations are into
``/root/reference``.  Parameters travel as a plain ``dict`` with the reference's
``state_dict`` keys (model.py:12-36): ``vis_layers.{0,2,4}.{weight,bias}``,
``lstm.{weight_ih,weight_hh,bias_ih,bias_hh}``, ``val.{0,2}.*``, ``adv.{0,2}.*``.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

PARAM_KEYS = (
    "vis_layers.0.weight", "vis_layers.0.bias",
    "vis_layers.2.weight", "vis_layers.2.bias",
    "vis_layers.4.weight", "vis_layers.4.bias",
    "lstm.weight_ih", "lstm.weight_hh", "lstm.bias_ih", "lstm.bias_hh",
    "val.0.weight", "val.0.bias", "val.2.weight", "val.2.bias",
    "adv.0.weight", "adv.0.bias", "adv.2.weight", "adv.2.bias",
)
N_ACTIONS = 6
FEAT = 7 * 7 * 32


def init_params(seed: int, cell: int = 256) -> dict:
    """Default-initialised parameters, drawn in the order ``QNet.__init__`` constructs its
    layers (model.py:12-36) so that ``torch.manual_seed(seed); QNet()`` gives the same
    numbers for cell == 256.  For other widths only the LSTM/head shapes change."""
    torch.manual_seed(seed)
    convs = [torch.nn.Conv2d(4, 32, 8, 4), torch.nn.Conv2d(32, 32, 4, 2),
             torch.nn.Conv2d(32, 32, 3, 1)]
    lstm = torch.nn.LSTMCell(FEAT, cell)
    val0, val2 = torch.nn.Linear(cell, cell), torch.nn.Linear(cell, 1)
    adv0, adv2 = torch.nn.Linear(cell, cell), torch.nn.Linear(cell, N_ACTIONS)
    p = {}
    for i, c in zip((0, 2, 4), convs):
        p[f"vis_layers.{i}.weight"] = c.weight.detach().clone()
        p[f"vis_layers.{i}.bias"] = c.bias.detach().clone()
    for k in ("weight_ih", "weight_hh", "bias_ih", "bias_hh"):
        p[f"lstm.{k}"] = getattr(lstm, k).detach().clone()
    for name, (l0, l2) in (("val", (val0, val2)), ("adv", (adv0, adv2))):
        p[f"{name}.0.weight"], p[f"{name}.0.bias"] = l0.weight.detach().clone(), l0.bias.detach().clone()
        p[f"{name}.2.weight"], p[f"{name}.2.bias"] = l2.weight.detach().clone(), l2.bias.detach().clone()
    return p


def torso(p: dict, frames: torch.Tensor, masks=None) -> torch.Tensor:
    """[N,4,84,84] f32 in [0,1] -> [N,1568], channel-major flatten (model.py:13-21,46).

    masks = (m1, m2, m3): 0/1 tensors shaped like the three conv outputs.  When given, each
    ReLU is evaluated as ``z * m`` -- the network with its activation pattern pinned.  Tests use
    it to compare gradients across two fp32 implementations whose pre-activations straddle a
    ReLU kink at a handful of positions (|z| ~ 1e-8): with the pattern of the implementation
    under test pinned, the remaining difference is pure rounding."""
    act = (lambda z, i: z * masks[i]) if masks is not None else (lambda z, i: F.relu(z))
    x = act(F.conv2d(frames, p["vis_layers.0.weight"], p["vis_layers.0.bias"], stride=4), 0)
    x = act(F.conv2d(x, p["vis_layers.2.weight"], p["vis_layers.2.bias"], stride=2), 1)
    x = act(F.conv2d(x, p["vis_layers.4.weight"], p["vis_layers.4.bias"], stride=1), 2)
    return x.reshape(x.shape[0], FEAT)


def torso_preacts(p: dict, frames: torch.Tensor):
    """(z1, z2, z3): the three conv pre-activations (no grad)."""
    with torch.no_grad():
        z1 = F.conv2d(frames, p["vis_layers.0.weight"], p["vis_layers.0.bias"], stride=4)
        z2 = F.conv2d(F.relu(z1), p["vis_layers.2.weight"], p["vis_layers.2.bias"], stride=2)
        z3 = F.conv2d(F.relu(z2), p["vis_layers.4.weight"], p["vis_layers.4.bias"], stride=1)
    return z1, z2, z3


def torso_kink_margin(p: dict, frames: torch.Tensor) -> float:
    """min |pre-activation| over the three conv layers: distance of the batch from a ReLU kink
    (tests use it to avoid comparing gradients across a discontinuity)."""
    with torch.no_grad():
        z1 = F.conv2d(frames, p["vis_layers.0.weight"], p["vis_layers.0.bias"], stride=4)
        z2 = F.conv2d(F.relu(z1), p["vis_layers.2.weight"], p["vis_layers.2.bias"], stride=2)
        z3 = F.conv2d(F.relu(z2), p["vis_layers.4.weight"], p["vis_layers.4.bias"], stride=1)
    return float(min(z1.abs().min(), z2.abs().min(), z3.abs().min()))


class _Bf16Linear(torch.autograd.Function):
    """x @ w.T with both operands rounded to bf16 and fp32 accumulation, in forward AND in both
    backward contractions -- the arithmetic of the tensor-core path (tc_gemm.cu), so that the
    bf16 kernels can be checked tightly against a same-precision restatement."""

    @staticmethod
    def forward(ctx, x, w):
        ctx.save_for_backward(x, w)
        return x.bfloat16().float() @ w.bfloat16().float().T

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        d = dy.bfloat16().float()
        return d @ w.bfloat16().float(), d.T @ x.bfloat16().float()


# which contractions are emulated in bf16 (set by train_step(..., tc_emulation=...))
_TC = {"proj": False, "head": False}


def _linear(x, w, b, kind):
    if _TC.get(kind):
        y = _Bf16Linear.apply(x, w)
        return y + b if b is not None else y
    return F.linear(x, w, b)


def lstm_step(p: dict, x, h, c):
    """One LSTMCell step, gate order i,f,g,o (model.py:24,55)."""
    g = _linear(x, p["lstm.weight_ih"], p["lstm.bias_ih"], "proj") + \
        F.linear(h, p["lstm.weight_hh"], p["lstm.bias_hh"])
    i, f, gg, o = g.chunk(4, dim=1)
    c2 = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(gg)
    h2 = torch.sigmoid(o) * torch.tanh(c2)
    return h2, c2


def head(p: dict, hs: torch.Tensor) -> torch.Tensor:
    """Dueling head: val + adv - mean_a(adv) (model.py:26-36,62-64)."""
    val = F.linear(F.relu(_linear(hs, p["val.0.weight"], p["val.0.bias"], "head")),
                   p["val.2.weight"], p["val.2.bias"])
    adv = F.linear(F.relu(_linear(hs, p["adv.0.weight"], p["adv.0.bias"], "head")),
                   p["adv.2.weight"], p["adv.2.bias"])
    return val + adv - adv.mean(1, keepdim=True)


def qnet_forward(p: dict, state: torch.Tensor, h, c, masks=None):
    """``QNet.forward`` with the module-held hidden state made explicit (model.py:38-72).

    state: [T,B,4,84,84] or [B,4,84,84] f32.  h,c: [B,H] or None (-> zeros, model.py:48-50).
    Returns q [T*B,6] (time-major rows t*B+b), hs_seq [T*B,H], cs_seq [T*B,H], (h,c) final.
    """
    if state.dim() == 5:
        T, B = state.shape[:2]
    else:
        T, B = 1, state.shape[0]
    H = p["lstm.weight_hh"].shape[1]
    feats = torso(p, state.reshape(-1, 4, 84, 84), masks).reshape(T, B, FEAT)
    if h is None:
        h = torch.zeros(B, H)
        c = torch.zeros(B, H)
    hs, cs = [], []
    for t in range(T):
        h, c = lstm_step(p, feats[t], h, c)
        hs.append(h)
        cs.append(c)
    hs_seq, cs_seq = torch.cat(hs, 0), torch.cat(cs, 0)
    return head(p, hs_seq), hs_seq, cs_seq, (h, c)


def value_rescale(x, eps=1e-3):
    """h(x) = sign(x)(sqrt(|x|+1)-1) + eps*x  (R2D2 paper; NOT in the reference -- the
    reference target is un-rescaled, learner.py:94.  Only used with rescale=True)."""
    return torch.sign(x) * (torch.sqrt(torch.abs(x) + 1.0) - 1.0) + eps * x


def value_rescale_inv(x, eps=1e-3):
    """h^-1(x) = sign(x)(((sqrt(1+4eps(|x|+1+eps))-1)/(2eps))^2 - 1)."""
    return torch.sign(x) * (((torch.sqrt(1.0 + 4.0 * eps * (torch.abs(x) + 1.0 + eps)) - 1.0)
                             / (2.0 * eps)) ** 2 - 1.0)


def rmsprop_centered(param, grad, square_avg, grad_avg, lr=0.00025 / 4.0, alpha=0.95,
                     eps=1.5e-7):
    """torch.optim.RMSprop(centered=True), no momentum/weight-decay (learner.py:51):
    v = a v + (1-a) g^2 ; m = a m + (1-a) g ; theta -= lr * g / (sqrt(v - m^2) + eps).
    Operates in place on f32 tensors, in torch's operation order."""
    square_avg.mul_(alpha).addcmul_(grad, grad, value=1 - alpha)
    grad_avg.lerp_(grad, 1 - alpha)
    avg = torch.addcmul(square_avg, grad_avg, grad_avg, value=-1).sqrt_().add_(eps)
    param.addcdiv_(grad, avg, value=-lr)


def train_step(p_online: dict, p_target: dict, batch: dict, *, burn_in: int, n_step: int,
               gamma: float = 0.99, alpha: float = 0.6, prio_eps: float = 1e-6,
               rescale: bool = False, next_action_override=None, tc_emulation=(),
               learn_masks=None) -> dict:
    """``Learner.train`` from the sampled batch to the per-step priorities
    (learner.py:71-102), without the optimiser step (see ``rmsprop_centered``).

    The online net's hidden state carries through its three passes (burn-in, learn,
    double-Q action pass) exactly as the reference's stateful module does.
    `batch` holds torch f32 tensors in the reference's layouts; `action` is int64.
    """
    for k in _TC:
        _TC[k] = k in tc_emulation
    try:
        return _train_step(p_online, p_target, batch, burn_in, n_step, gamma, alpha, prio_eps,
                           rescale, next_action_override, learn_masks)
    finally:
        for k in _TC:
            _TC[k] = False


def _train_step(p_online, p_target, batch, burn_in, n_step, gamma, alpha, prio_eps, rescale,
                next_action_override, learn_masks=None):
    po = {k: v.detach().clone().requires_grad_(True) for k, v in p_online.items()}
    BI = burn_in
    state, next_state = batch["state"], batch["next_state"]
    h, c = batch["hs"], batch["cs"]
    th, tc = batch["target_hs"], batch["target_cs"]
    with torch.no_grad():
        _, _, _, (h, c) = qnet_forward(po, state[:BI], h, c)
        _, _, _, (th, tc) = qnet_forward(p_target, next_state[:BI], th, tc)
    # learn_masks: ReLU pattern of the grad-enabled pass pinned (see torso)
    q_all, hs_seq, cs_seq, (h, c) = qnet_forward(po, state[BI:], h, c, learn_masks)
    q = q_all.gather(1, batch["action"].view(-1, 1))
    ends = {"h_learn_end": h.detach(), "c_learn_end": c.detach()}
    with torch.no_grad():
        q_next_online, _, _, (hd, cd) = qnet_forward(po, next_state[BI:], h.detach(), c.detach())
        ends["h_dq_end"], ends["c_dq_end"] = hd, cd
        next_action = torch.argmax(q_next_online, 1).view(-1, 1)
        if next_action_override is not None:
            next_action = next_action_override.view(-1, 1)
        q_next_target, _, _, (th, tc) = qnet_forward(p_target, next_state[BI:], th, tc)
        ends["h_tg_end"], ends["c_tg_end"] = th, tc
        nq = q_next_target.gather(1, next_action)
        r = batch["reward"].view(-1, 1)
        d = batch["done"].view(-1, 1)
        if rescale:
            y = value_rescale(r + (gamma ** n_step) * value_rescale_inv(nq) * (1 - d))
        else:
            y = r + (gamma ** n_step) * nq * (1 - d)
    loss = torch.mean(0.5 * (q - y) ** 2)
    grads = torch.autograd.grad(loss, [po[k] for k in PARAM_KEYS])
    delta = (q - y).detach()
    prio = (np.abs(delta.numpy()).reshape(-1) + prio_eps) ** alpha
    return dict(q_all=q_all.detach(), q=q.detach(), q_next_online=q_next_online,
                q_next_target=q_next_target, next_action=next_action.view(-1),
                y=y, delta=delta, loss=loss.detach(), priority=prio.astype(np.float32),
                grads={k: g for k, g in zip(PARAM_KEYS, grads)},
                hs_seq=hs_seq.detach(), cs_seq=cs_seq.detach(), **ends)


def to_torch_batch(b: dict) -> dict:
    """numpy batch (``replay_oracle.gather_batch``) -> torch, as replay_memory.py:258-260."""
    out = {k: torch.from_numpy(np.ascontiguousarray(v)).float() for k, v in b.items()}
    out["action"] = out["action"].long()
    return out
