"""GPU-resident prioritized sequence replay behind the reference's ``ReplayMemory`` API.

Mirrors /root/reference replay_memory.py:58-277: same constructor, attributes, the ten
``memory[...]`` keys with the same dtypes and shapes, and the methods ``add``, ``extend``,
``sample``, ``update_priority``, ``update_sequence_priority`` (plus the unused
``set_hs_cs``/``get_stacked_state``/``indexing_sample``/``fit``/``save``/``load``).  Storage lives in
HBM; index selection is a 32-ary f64 sum-tree descent (K1) instead of the O(M) mask +
``torch.multinomial``; the slab gather, priority scatter, sequence-priority refresh and tree
update are CUDA kernels called through the C ABI.  There is no CPU path.

North-star aliases: ``push`` = ``add``; ``update_priorities(index, priority, seq_index)`` =
``update_priority`` + ``update_sequence_priority(seq_index, True)``.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

from . import _lib

KEYS = ("state", "hs_cs", "target_hs_cs", "action", "reward", "done", "stack_count",
        "priority", "sequence_priority", "is_seq_start")
_TREE_KEYS = ("sequence_priority", "is_seq_start")


class _Memory(dict):
    """``replay.memory``: the reference lets callers poke arrays directly
    (actor.py:167 sets ``memory['is_seq_start'][idx] = 1``).  Handing out a tensor that feeds
    the sum-tree marks the tree stale; it is rebuilt before the next draw."""

    def __init__(self, owner):
        super().__init__()
        self._owner = owner

    def __getitem__(self, key):
        if key in _TREE_KEYS:
            self._owner._tree_dirty = True
        return super().__getitem__(key)


def tree_layout(M: int):
    sizes = (ctypes.c_int64 * 8)()
    offs = (ctypes.c_int64 * 8)()
    nlev = ctypes.c_int(0)
    total = ctypes.c_int64(0)
    _lib.call("r2d2_tree_layout", M, 8, sizes, offs, ctypes.byref(nlev), ctypes.byref(total))
    return list(sizes[: nlev.value]), list(offs[: nlev.value]), total.value


class ReplayMemory:
    def __init__(self, memory_size=100000, batch_size=32, n_step=3, state_size=(84, 84),
                 cell_size=256, action_repeat=4, n_stacks=4, alpha=0.4, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("pytorch_r2d2_b200.ReplayMemory needs a CUDA device (no CPU fallback)")
        _lib.load()
        if tuple(state_size) != (84, 84) or action_repeat != 4:
            raise ValueError("the CUDA replay stores 4x84x84 u8 frame blocks (replay_memory.py:65,77)")
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.index = 0
        self.memory_size = memory_size
        self.cell_size = cell_size
        self.batch_size = batch_size
        self.n_step = n_step
        self.state_size = (action_repeat,) + tuple(state_size)
        self.action_repeat = action_repeat
        self.n_stacks = n_stacks // action_repeat
        self.alpha = alpha
        self.beta = 0.4
        self.beta_step = 0.00025 / 4
        self.eta = 0.9
        self.burn_in_length = 10
        self.learning_length = 10
        self.sequence_length = self.burn_in_length + self.learning_length
        self.use_tma_gather = False

        M, dev = memory_size, self.device
        mem = _Memory(self)
        dict.__setitem__(mem, "state", torch.zeros((M, *self.state_size), dtype=torch.uint8, device=dev))
        dict.__setitem__(mem, "hs_cs", torch.zeros((M, cell_size * 2), dtype=torch.float32, device=dev))
        dict.__setitem__(mem, "target_hs_cs", torch.zeros((M, cell_size * 2), dtype=torch.float32, device=dev))
        dict.__setitem__(mem, "action", torch.zeros((M, 1), dtype=torch.int8, device=dev))
        dict.__setitem__(mem, "reward", torch.zeros((M, 1), dtype=torch.float32, device=dev))
        dict.__setitem__(mem, "done", torch.zeros((M, 1), dtype=torch.float32, device=dev))
        dict.__setitem__(mem, "stack_count", torch.zeros((M,), dtype=torch.int8, device=dev))
        dict.__setitem__(mem, "priority", torch.zeros((M,), dtype=torch.float32, device=dev))
        dict.__setitem__(mem, "sequence_priority", torch.zeros((M,), dtype=torch.float32, device=dev))
        dict.__setitem__(mem, "is_seq_start", torch.zeros((M,), dtype=torch.uint8, device=dev))
        self.memory = mem
        self.arange = torch.arange(M, device=dev)

        _, _, total = tree_layout(M)
        self._nodes = torch.zeros(total, dtype=torch.float64, device=dev)
        self._tree_dirty = True
        self._ingest_stream, self._pending, self._stage, self._stage_turn = None, [], {}, 0
        self._u_ring, self._u_turn = None, 0   # pinned staging slots for the uniforms
        self._ws = {}            # sample workspaces keyed by shape
        self._touched = None

    # ------------------------------------------------------------------ helpers
    def _raw(self, key):
        return dict.__getitem__(self.memory, key)

    @property
    def size(self):
        return min(self.index, self.memory_size)

    def _ensure_tree(self):
        self._wait_ingest()
        if self._tree_dirty:
            _lib.call("r2d2_tree_build", _lib.ptr(self._raw("sequence_priority"), "f32"),
                      _lib.ptr(self._raw("is_seq_start"), "u8"), self.memory_size,
                      _lib.ptr(self._nodes, "f64"), _lib.stream_ptr())
            self._tree_dirty = False

    def rebuild_tree(self):
        self._tree_dirty = True
        self._ensure_tree()

    def _dev(self, x, dtype):
        t = torch.as_tensor(np.asarray(x) if not torch.is_tensor(x) else x)
        return t.to(device=self.device, dtype=dtype, non_blocking=True).contiguous()

    # ------------------------------------------------------------------ write path
    _DT = {"state": "u8", "hs_cs": "f32", "target_hs_cs": "f32", "action": "i8", "reward": "f32",
           "done": "f32", "stack_count": "i8", "priority": "f32", "sequence_priority": "f32",
           "is_seq_start": "u8"}

    def _staging(self, rows, slot):
        """Persistent device staging buffers (two sets, ping-pong) for overlapped ingest."""
        key = (rows, slot)
        st = self._stage.get(key)
        if st is None:
            H, dev = self.cell_size, self.device
            st = {
                "state": torch.empty((rows, 28224), dtype=torch.uint8, device=dev),
                "hs_cs": torch.empty((rows, 2 * H), device=dev),
                "target_hs_cs": torch.empty((rows, 2 * H), device=dev),
                "action": torch.empty(rows, dtype=torch.int8, device=dev),
                "reward": torch.empty(rows, device=dev), "done": torch.empty(rows, device=dev),
                "stack_count": torch.empty(rows, dtype=torch.int8, device=dev),
                "priority": torch.empty(rows, device=dev),
                "sequence_priority": torch.empty(rows, device=dev),
                "is_seq_start": torch.empty(rows, dtype=torch.uint8, device=dev),
                "free": torch.cuda.Event(), "ready": torch.cuda.Event(),
            }
            st["free"].record(torch.cuda.current_stream())
            self._stage[key] = st
        return st

    def extend(self, memory, overlap=False):
        """Bulk-append an actor dump (replay_memory.py:107-119): all ten keys, ring wrap.

        overlap=True splits the append in two: the host->device copies go to a private staging
        buffer on a dedicated ingest stream (they overlap whatever the learner is running and
        touch nothing it reads); the scatter from staging into the ring (an HBM-speed kernel) is
        enqueued on the caller's stream right before the next draw.  A draw therefore always
        sees every dump appended before it, and rows are never overwritten under a running
        gather."""
        rows = int(memory["state"].shape[0])
        if rows == 0:
            return
        if rows > self.memory_size:
            raise ValueError("dump larger than the ring")
        if not overlap:
            self._wait_ingest()
            src = {k: self._dev(memory[k], _lib._DT[self._DT[k]]).reshape(rows, -1) if k in
                   ("state", "hs_cs", "target_hs_cs") else self._dev(memory[k], _lib._DT[self._DT[k]]).reshape(rows)
                   for k in KEYS}
            self._scatter(src, rows, self.index % self.memory_size)
        else:
            if self._ingest_stream is None:
                self._ingest_stream = torch.cuda.Stream(device=self.device)
            st = self._staging(rows, self._stage_turn)
            self._stage_turn ^= 1
            if any(p[0] is st for p in self._pending):
                # a third dump of this size before the next draw: the set still holds rows whose
                # scatter has not been enqueued -- land everything staged so far first
                self._wait_ingest()
            with torch.cuda.stream(self._ingest_stream):
                self._ingest_stream.wait_event(st["free"])     # previous scatter out of this set done
                for k in KEYS:
                    srcv = memory[k] if torch.is_tensor(memory[k]) else torch.as_tensor(np.asarray(memory[k]))
                    st[k].copy_(srcv.reshape(st[k].shape), non_blocking=True)
                st["ready"].record(self._ingest_stream)
            self._pending.append((st, rows, self.index % self.memory_size))
        self.index += rows
        self._tree_dirty = True

    def _scatter(self, src, rows, write_index):
        args = [_lib.ptr(src[k], self._DT[k]) for k in KEYS] + [rows] + \
               [_lib.ptr(self._raw(k), self._DT[k]) for k in KEYS] + \
               [self.memory_size, self.cell_size, write_index, _lib.stream_ptr()]
        _lib.call("r2d2_replay_ingest", *args)

    def _wait_ingest(self):
        """Land every staged dump in the ring (on the current stream, in append order)."""
        if self._pending:
            cur = torch.cuda.current_stream()
            for st, rows, widx in self._pending:
                cur.wait_event(st["ready"])
                self._scatter(st, rows, widx)
                st["free"].record(cur)
            self._pending = []

    def add(self, state, hs, cs, target_hs, target_cs, action, reward, done, stack_count, priority):
        """One transition (replay_memory.py:93-105).  `state` is a float frame block in [0,1];
        it is stored as ``state*255`` truncated to u8.  Like the reference, this does not touch
        ``sequence_priority`` / ``is_seq_start`` of the slot."""
        i = self.index % self.memory_size
        H = self.cell_size
        frame = (torch.as_tensor(np.asarray(state, dtype=np.float32)) * 255).to(torch.uint8)
        m = self.memory
        self._raw("state")[i].copy_(frame.reshape(self.state_size), non_blocking=True)
        hc = torch.as_tensor(np.concatenate([np.asarray(hs, np.float32).reshape(-1),
                                             np.asarray(cs, np.float32).reshape(-1)]))
        thc = torch.as_tensor(np.concatenate([np.asarray(target_hs, np.float32).reshape(-1),
                                              np.asarray(target_cs, np.float32).reshape(-1)]))
        assert hc.numel() == 2 * H
        self._raw("hs_cs")[i].copy_(hc, non_blocking=True)
        self._raw("target_hs_cs")[i].copy_(thc, non_blocking=True)
        self._raw("action")[i] = int(action)
        self._raw("reward")[i] = float(reward)
        self._raw("done")[i] = 1.0 if done else 0.0
        self._raw("stack_count")[i] = int(stack_count)
        self._raw("priority")[i] = float(priority)
        self.index += 1
        del m

    push = add

    def fit(self):
        """The reference truncates its arrays to ``size`` before a dump (replay_memory.py:121-123).
        The device ring keeps its capacity; ``dump()`` applies the truncation to what it returns."""
        return None

    def dump(self):
        """``{key: ndarray[:size]}`` with the reference's ten keys, dtypes and shapes: the payload
        of the file hand-off and of ``extend``."""
        n = self.size
        self._wait_ingest()
        return {k: self._raw(k)[:n].cpu().numpy() for k in KEYS}

    @staticmethod
    def _lock_file(path, timeout):
        """The reference guards each dump file with ``fasteners.InterProcessLock(path)``
        (replay_memory.py:127,157): an fcntl lock on the DATA FILE itself, opened 'a+' (which
        creates it empty -- hence the size checks), polled without blocking.  Same protocol here
        (``fcntl.lockf``, what fasteners uses on POSIX), so this learner and the reference's
        actors can share a directory.  Returns the locked file object, or None on time-out.
        A file that was removed and recreated while we waited is a different inode: retry."""
        import fcntl
        import time
        deadline = time.monotonic() + timeout
        while True:
            f = open(path, "a+b")
            try:
                fcntl.lockf(f, fcntl.LOCK_EX | fcntl.LOCK_NB)
                try:
                    if os.fstat(f.fileno()).st_ino == os.stat(path).st_ino:
                        return f
                except FileNotFoundError:
                    pass
                fcntl.lockf(f, fcntl.LOCK_UN)
            except OSError:
                pass
            f.close()
            if time.monotonic() >= deadline:
                return None
            time.sleep(0.05)

    def save(self, path, actor_id, timeout=30.0):
        """File hand-off, writer side (replay_memory.py:125-152): under the per-file lock, merge
        with a dump that is still waiting there (own rows first, then the waiting ones, as
        ``extend`` + ``fit`` give), and write ``torch.save({key: ndarray[:size]})``.  The bytes are
        serialised in memory first and written in place under the lock, so a reader that honours
        the lock never sees a torn file.  An unreadable waiting dump is replaced, not merged."""
        import io
        os.makedirs(path, exist_ok=True)
        path = os.path.join(path, f"memory{actor_id}.pt")
        new = self.dump()
        f = self._lock_file(path, timeout)
        if f is None:
            raise TimeoutError(f"could not lock {path}")
        try:
            f.seek(0)
            raw = f.read()
            if raw:
                try:
                    old = torch.load(io.BytesIO(raw), weights_only=False)
                    new = {k: np.concatenate([new[k], old[k]]) for k in KEYS}
                except Exception:      # noqa: BLE001 -- a corrupt leftover must not block the hand-off
                    pass
            buf = io.BytesIO()
            torch.save(new, buf)
            f.seek(0)
            f.truncate()
            f.write(buf.getvalue())
            f.flush()
            os.fsync(f.fileno())
        finally:
            f.close()                  # closing releases the lock

    def load(self, path, actor_id, timeout=0.0):
        """File hand-off, reader side (replay_memory.py:155-173): if a non-empty dump is waiting
        and its lock is free, ``extend`` with it and delete the file.  Like the reference a busy
        lock means "try again later" (returns False).  Unlike the reference an unreadable file is
        NOT deleted: the error is raised and the file stays for inspection."""
        import io
        path = os.path.join(path, f"memory{actor_id}.pt")
        if not (os.path.isfile(path) and os.path.getsize(path) > 0):
            return False
        f = self._lock_file(path, timeout)
        if f is None:
            return False
        try:
            f.seek(0)
            raw = f.read()
            if not raw:
                return False
            memory = torch.load(io.BytesIO(raw), weights_only=False)
            self.extend(memory)
            os.remove(path)            # only after a successful read, still under the lock
            return True
        finally:
            f.close()

    # ------------------------------------------------------------------ priorities
    def update_priority(self, index, priority):
        """priority[index] = priority, last write wins on duplicates (replay_memory.py:175-176)."""
        idx = self._dev(index, torch.int64).reshape(-1)
        p = self._dev(priority, torch.float32).reshape(-1)
        if idx.numel() != p.numel():
            raise ValueError("index and priority must have the same number of elements")
        _lib.call("r2d2_update_priority", _lib.ptr(self._raw("priority"), "f32"), self.memory_size,
                  _lib.ptr(idx, "i64"), _lib.ptr(p, "f32"), idx.numel(), _lib.stream_ptr())

    def update_sequence_priority(self, index, update_pre_next_seq_priority=False):
        """replay_memory.py:184-213 including its neighbour quirks; then refresh the tree."""
        idx = self._dev(index, torch.int64).reshape(-1)
        n = idx.numel()
        if n == 0:
            return
        if self._touched is None or self._touched.numel() < 3 * n:
            self._touched = torch.empty(3 * n, dtype=torch.int64, device=self.device)
        eta32 = float(np.float32(self.eta))
        ometa32 = float(np.float32(1 - self.eta))
        _lib.call("r2d2_update_sequence_priority", _lib.ptr(self._raw("priority"), "f32"),
                  _lib.ptr(self._raw("sequence_priority"), "f32"),
                  _lib.ptr(self._raw("is_seq_start"), "u8"), self.memory_size, _lib.ptr(idx, "i64"),
                  n, self.sequence_length, eta32, ometa32, int(bool(update_pre_next_seq_priority)),
                  _lib.ptr(self._touched, "i64"), _lib.stream_ptr())
        if not self._tree_dirty:
            _lib.call("r2d2_tree_update", _lib.ptr(self._raw("sequence_priority"), "f32"),
                      _lib.ptr(self._raw("is_seq_start"), "u8"), self.memory_size,
                      _lib.ptr(self._nodes, "f64"), _lib.ptr(self._touched, "i64"), 3 * n,
                      _lib.stream_ptr())

    def update_priorities(self, index, priority, seq_index):
        self.update_priority(index, priority)
        self.update_sequence_priority(seq_index, True)

    def update_priority_learn_window(self, seq_index_dev, prio_dev):
        """Learner fast path of learner.py:103: scatter the [L,B] per-step priorities of the
        learn window straight from device memory."""
        B = seq_index_dev.numel()
        _lib.call("r2d2_update_priority_seq", _lib.ptr(self._raw("priority"), "f32"),
                  self.memory_size, _lib.ptr(seq_index_dev, "i64"), B, self.burn_in_length,
                  self.learning_length, _lib.ptr(prio_dev, "f32"), _lib.stream_ptr())

    def set_hs_cs(self, index, hs, cs, target_hs, target_cs):
        H = self.cell_size
        idx = self._dev(index, torch.int64).reshape(-1)
        self._raw("hs_cs")[idx, :H] = self._dev(hs, torch.float32)
        self._raw("hs_cs")[idx, H:] = self._dev(cs, torch.float32)
        self._raw("target_hs_cs")[idx, :H] = self._dev(target_hs, torch.float32)
        self._raw("target_hs_cs")[idx, H:] = self._dev(target_cs, torch.float32)

    # ------------------------------------------------------------------ sampling
    U_SLOTS = 8

    def upload_uniforms(self, B, u_dev):
        """Draw the B float64 uniforms ``torch.multinomial`` would consume from the global CPU
        generator (replay_memory.py:227-231) and enqueue their copy into `u_dev` on the current
        stream.  The host may run many steps ahead of the device (CUDA-graph replays cost it
        ~30 us against ~2 ms of device work), so the uniforms of every step get a pinned slot
        of their own out of a ring of U_SLOTS, and a slot is only redrawn once the copy that
        read it has completed (event recorded behind the copy; the wait only blocks when the
        host is more than U_SLOTS steps ahead).  Returns the pinned slot (valid until the ring
        wraps)."""
        if self._u_ring is None or self._u_ring[0][0].numel() != B:
            self._u_ring = [[torch.empty(B, dtype=torch.float64).pin_memory(), None]
                            for _ in range(self.U_SLOTS)]
        slot = self._u_ring[self._u_turn % self.U_SLOTS]
        self._u_turn += 1
        if slot[1] is None:
            slot[1] = torch.cuda.Event()
        else:
            slot[1].synchronize()
        torch.rand(B, dtype=torch.float64, out=slot[0])
        u_dev.copy_(slot[0], non_blocking=True)
        slot[1].record(torch.cuda.current_stream())
        return slot[0]

    def configure(self, batch_size, n_step, burn_in_length, learning_length):
        """Set the sampling shape (plain attributes in the reference: replay_memory.py:63-74)."""
        self.batch_size, self.n_step = batch_size, n_step
        self.burn_in_length, self.learning_length = burn_in_length, learning_length
        self.sequence_length = burn_in_length + learning_length

    def _workspace(self, B, T, n, H):
        key = (B, T, n, H, self.burn_in_length)
        ws = self._ws.get(key)
        if ws is None:
            dev, L = self.device, T - self.burn_in_length
            ws = dict(
                u=torch.empty(B, dtype=torch.float64, device=dev),
                seq_index=torch.empty(B, dtype=torch.int64, device=dev),
                slab=torch.empty((B, T + n, *self.state_size), dtype=torch.uint8, device=dev),
                hs=torch.empty((B, H), device=dev), cs=torch.empty((B, H), device=dev),
                target_hs=torch.empty((B, H), device=dev), target_cs=torch.empty((B, H), device=dev),
                action=torch.empty((L, B), dtype=torch.int64, device=dev),
                reward=torch.empty((L, B), device=dev), done=torch.empty((L, B), device=dev),
                index=torch.empty((T, B), dtype=torch.int64, device=dev))
            self._ws = {key: ws}
        return ws

    def sample_indices(self, u):
        """Sum-tree draw for the given float64 uniforms (device or host tensor)."""
        u = self._dev(u, torch.float64).reshape(-1)
        out = torch.empty(u.numel(), dtype=torch.int64, device=self.device)
        self._ensure_tree()
        _lib.call("r2d2_tree_sample", _lib.ptr(self._raw("sequence_priority"), "f32"),
                  _lib.ptr(self._raw("is_seq_start"), "u8"), self.memory_size,
                  _lib.ptr(self._nodes, "f64"), _lib.ptr(u, "f64"), u.numel(), _lib.ptr(out, "i64"),
                  _lib.stream_ptr())
        return out

    def sample_slab(self, u=None, seq_index=None, frames=True, hidden_out=None):
        """Device-resident fast path: draw, then gather one contiguous (T+n)-frame u8 slab per
        sequence plus hidden/action/reward/done/index.  Returns the workspace dict (tensors are
        reused by the next call).  frames=False leaves the frames in the ring (``ws["ring"]``,
        read in place through ``ws["seq_index"]`` by the ring forms of the first conv layer).
        hidden_out = (hs, cs, target_hs, target_cs) [B,H] f32 tensors that receive the initial
        recurrent states directly (e.g. slot 0 of the caller's history buffers)."""
        B, T, n, H = self.batch_size, self.sequence_length, self.n_step, self.cell_size
        ws = self._workspace(B, T, n, H)
        self._wait_ingest()
        if seq_index is None:
            if u is None:
                self.upload_uniforms(B, ws["u"])
            elif not (torch.is_tensor(u) and u.is_cuda and u.data_ptr() == ws["u"].data_ptr()):
                # caller-supplied uniforms (a device tensor that IS the workspace buffer was
                # uploaded by the caller already: the graphed learner step does that)
                ws["u"].copy_(u if torch.is_tensor(u) else torch.as_tensor(u), non_blocking=True)
            self._ensure_tree()
            _lib.call("r2d2_tree_sample", _lib.ptr(self._raw("sequence_priority"), "f32"),
                      _lib.ptr(self._raw("is_seq_start"), "u8"), self.memory_size,
                      _lib.ptr(self._nodes, "f64"), _lib.ptr(ws["u"], "f64"), B,
                      _lib.ptr(ws["seq_index"], "i64"), _lib.stream_ptr())
        else:
            ws["seq_index"].copy_(self._dev(seq_index, torch.int64).reshape(-1))
        ho = hidden_out if hidden_out is not None else (ws["hs"], ws["cs"], ws["target_hs"], ws["target_cs"])
        _lib.call("r2d2_replay_gather", _lib.ptr(self._raw("state"), "u8"),
                  _lib.ptr(self._raw("hs_cs"), "f32"), _lib.ptr(self._raw("target_hs_cs"), "f32"),
                  _lib.ptr(self._raw("action"), "i8"), _lib.ptr(self._raw("reward"), "f32"),
                  _lib.ptr(self._raw("done"), "f32"), self.memory_size, H,
                  _lib.ptr(ws["seq_index"], "i64"), B, T, n, self.burn_in_length,
                  _lib.ptr(ws["slab"], "u8") if frames else None, _lib.ptr(ho[0], "f32"), _lib.ptr(ho[1], "f32"),
                  _lib.ptr(ho[2], "f32"), _lib.ptr(ho[3], "f32"),
                  _lib.ptr(ws["action"], "i64"), _lib.ptr(ws["reward"], "f32"),
                  _lib.ptr(ws["done"], "f32"), _lib.ptr(ws["index"], "i64"),
                  int(self.use_tma_gather), _lib.stream_ptr())
        ws["ring"] = self._raw("state")
        return ws

    def sample(self, device=None):
        """Reference-compatible ``sample`` (replay_memory.py:224-262): returns
        ``(batch, seq_index, index)`` with f32 ``state``/``next_state`` [T,B,4,84,84] (u8/255),
        hidden states, ``action`` int64 / ``reward`` / ``done`` [L,B,1]; ``seq_index`` and
        ``index`` are numpy arrays like the reference's.  ``device`` is accepted for signature
        compatibility; tensors stay on this replay's CUDA device."""
        ws = self.sample_slab()
        B, T, n = self.batch_size, self.sequence_length, self.n_step
        S = T + n
        batch = {}
        for key, s0 in (("state", 0), ("next_state", n)):
            out = torch.empty((T, B, *self.state_size), dtype=torch.float32, device=self.device)
            _lib.call("r2d2_slab_to_f32", _lib.ptr(ws["slab"], "u8"), B, S, s0, T,
                      _lib.ptr(out, "f32"), _lib.stream_ptr())
            batch[key] = out
        for k in ("hs", "cs", "target_hs", "target_cs"):
            batch[k] = ws[k].clone()
        batch["action"] = ws["action"].clone().unsqueeze(-1)
        batch["reward"] = ws["reward"].clone().unsqueeze(-1)
        batch["done"] = ws["done"].clone().unsqueeze(-1)
        return batch, ws["seq_index"].cpu().numpy(), ws["index"].cpu().numpy()

    def get_stacked_state(self, index):
        """replay_memory.py:215-222 (general stacking rule), evaluated on the host copy of the
        few rows involved; the gather kernel implements the stack_count == n_stacks case."""
        sc = int(self._raw("stack_count")[index].item())
        start = index - (self.n_stacks - sc)
        if start < 0:
            start += self.memory_size
        rows = [start] * sc + [(start + 1 + i) % self.memory_size for i in range(self.n_stacks - sc)]
        return torch.cat([self._raw("state")[i] for i in rows]).cpu().numpy()

    def indexing_sample(self, start_index, last_index, device=None):
        """replay_memory.py:264-277."""
        M = self.memory_size
        index = np.arange(start_index, last_index) % M
        nxt = (index + self.n_step) % M
        batch = {
            "state": np.stack([[self.get_stacked_state(int(i))] for i in index]).astype(np.float32) / 255.,
            "next_state": np.stack([[self.get_stacked_state(int(i))] for i in nxt]).astype(np.float32) / 255.,
            "action": self._raw("action")[index].cpu().numpy(),
            "reward": self._raw("reward")[index].cpu().numpy(),
            "done": self._raw("done")[index].cpu().numpy(),
        }
        return batch, index
