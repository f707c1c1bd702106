"""Drive the UNMODIFIED reference (byte-compiled into ``oracle/_ref`` by ``oracle/make_ref.py``)
at arbitrary shapes: the CPU baseline of ``bench.py`` and, with ``device='cuda:0'``, the
informational "PyTorch eager on the same B200" line (the reference's own default device,
learner.py:19).

TEST / BASELINE INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.

What has to be set from outside, because the reference hard-codes it:
lengths / batch / n are plain attributes (learner.py:23,35-39, replay_memory.py:63-74); the LSTM
width 256 is a literal in ``QNet.__init__`` (model.py:24-35), so for the 512-wide configuration
the ``lstm`` / ``val`` / ``adv`` sub-modules are replaced by wider ones of the same classes --
``forward``'s body runs unchanged (its own 256 only appears in the zero-state branch, which
``train()`` never takes: ``set_state`` precedes every forward, learner.py:71-72).
``fasteners`` (file hand-off lock, not on the hot path, not installed) is stubbed.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")


def available() -> bool:
    return all(os.path.exists(os.path.join(REF_DIR, m + ".ref")) for m in ("replay_memory", "model", "learner"))


def load_reference():
    """(replay_memory, model, learner) modules of the reference, or None if oracle/_ref is absent."""
    if not available():
        return None
    if "fasteners" not in sys.modules:
        sys.modules["fasteners"] = types.ModuleType("fasteners")
    import importlib.machinery
    import importlib.util
    mods = []
    # dependency order: learner.py does `from model import QNet` / `from replay_memory import ...`
    for name in ("replay_memory", "model", "learner"):
        path = os.path.join(REF_DIR, name + ".ref")
        m = sys.modules.get(name)
        if m is None or getattr(m, "__file__", None) != path:
            loader = importlib.machinery.SourcelessFileLoader(name, path)
            spec = importlib.util.spec_from_loader(name, loader, origin=path)
            m = importlib.util.module_from_spec(spec)
            m.__file__ = path
            sys.modules[name] = m
            loader.exec_module(m)
        mods.append(m)
    return tuple(mods)


def widen(net, H, device):
    """Replace the 256-wide sub-modules of a reference QNet by H-wide ones (same classes)."""
    if H == 256:
        return net
    net.lstm = nn.LSTMCell(7 * 7 * 32, H)
    net.val = nn.Sequential(nn.Linear(H, H), nn.ReLU(True), nn.Linear(H, 1))
    net.adv = nn.Sequential(nn.Linear(H, H), nn.ReLU(True), nn.Linear(H, 6))
    return net.to(device)


def fill_tiled(mem: dict, N: int, H: int, seed=1, block=20000, done_period=997):
    """SYN-R2D2 contents for rings too large to draw element by element:
    a `block`-row random block repeated over the first N rows."""
    rng = np.random.RandomState(seed)
    nb = min(block, N)
    blk = {
        "state": rng.randint(0, 256, (nb, 4, 84, 84)).astype(np.uint8),
        "hs_cs": (rng.randn(nb, 2 * H) * 0.1).astype(np.float32),
        "target_hs_cs": (rng.randn(nb, 2 * H) * 0.1).astype(np.float32),
        "action": rng.randint(6, size=(nb, 1)).astype(np.int8),
        "reward": rng.randn(nb, 1).astype(np.float32),
        "priority": (rng.rand(nb) + 0.01).astype(np.float32),
    }
    for lo in range(0, N, nb):
        n = min(nb, N - lo)
        for k, v in blk.items():
            mem[k][lo:lo + n] = v[:n]
    mem["done"][:N, 0] = (np.arange(N) % done_period == done_period - 1).astype(np.float32)
    mem["stack_count"][:N] = 1


def make_reference_learner(B, BI, L, n, H, M, device="cpu", seed=0, stride=40):
    """The reference's ``Learner`` on `device` at the given shape, its replay filled with
    SYN-R2D2 (80 % of M slots, a sequence start every `stride` slots)."""
    mods = load_reference()
    if mods is None:
        raise RuntimeError("oracle/_ref is missing: run `python oracle/make_ref.py` in the build container")
    ref_replay, ref_model, ref_learner = mods
    torch.manual_seed(seed)
    # the constructor allocates its default 500 000-slot ring with np.zeros (untouched pages cost
    # nothing) and is replaced right away
    Lr = ref_learner.Learner(1, {}, device=device)
    Lr.net = widen(Lr.net, H, device)
    Lr.target_net = widen(Lr.target_net, H, device)
    Lr.target_net.load_state_dict(Lr.net.state_dict())
    Lr.optim = torch.optim.RMSprop(Lr.net.parameters(), lr=0.00025 / 4.0, alpha=0.95, eps=1.5e-7, centered=True)
    rm = ref_replay.ReplayMemory(M, B, n, cell_size=H)
    for obj in (Lr, rm):
        obj.burn_in_length, obj.learning_length, obj.sequence_length = BI, L, BI + L
    Lr.bootstrap_steps, Lr.batch_size, Lr.memory_size = n, B, M
    N = int(0.8 * M)
    fill_tiled(rm.memory, N, H)
    rm.index = N
    starts = np.arange(0, max(N - (BI + L) - n, 1), stride)
    for i in range(0, len(starts), 4096):
        rm.update_sequence_priority(starts[i:i + 4096])
    rm.memory["is_seq_start"][starts] = 1
    Lr.replay_memory = rm
    return Lr
