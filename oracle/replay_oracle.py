"""Oracle (CPU restatement) of the reference's prioritized sequence replay.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.  numpy, f32/f64 arithmetic in the
exact order the reference (numpy 2.x / torch 2.11 CPU) performs it.  Citations are into
``/root/reference``.

Contents
  * ``np_pairwise_sum_f32``      -- numpy's float32 ``add.reduce`` order (used by ``mean``),
                                    restated so the CUDA kernel can mirror it bit for bit.
  * ``seq_priority_value``       -- eta*max + (1-eta)*mean       (replay_memory.py:187-188)
  * ``update_priority``          -- priority[index] = p          (replay_memory.py:175-176)
  * ``update_sequence_priority`` -- incl. the neighbour quirks   (replay_memory.py:184-213)
  * ``ref_sample_indices``       -- inverse-CDF semantics of WeightedRandomSampler /
                                    torch.multinomial            (replay_memory.py:225-233)
  * ``SumTree``                  -- the 32-ary f64 tree the CUDA path uses (no reference
                                    counterpart; contract tiers in DESIGN.md)
  * ``gather_batch``             -- the slab gather               (replay_memory.py:235-262)
  * ``ReplayOracle``             -- the ring buffer with add/extend (replay_memory.py:59-119)
"""
from __future__ import annotations

import numpy as np

FAN = 32  # sum-tree fan-out == warp width


# --------------------------------------------------------------------------------------
# numpy float32 reduction order
# --------------------------------------------------------------------------------------
def np_pairwise_sum_f32(a: np.ndarray) -> np.float32:
    """Sum of a contiguous float32 vector in the order numpy's ``add.reduce`` uses.

    Mirrors numpy's pairwise summation (used by ``ndarray.sum`` / ``mean``, which the
    reference calls at replay_memory.py:188,202,213): n < 8 -> plain left-to-right loop;
    n <= 128 -> eight strided partial sums combined as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)),
    tail added left to right; larger n -> split at n/2 rounded down to a multiple of 8.
    """
    a = np.ascontiguousarray(a, dtype=np.float32)
    n = a.shape[0]
    f = np.float32
    if n < 8:
        res = f(0.0)
        for i in range(n):
            res = f(res + a[i])
        return res
    if n <= 128:
        r = [f(a[j]) for j in range(8)]
        i = 8
        while i < n - (n % 8):
            for j in range(8):
                r[j] = f(r[j] + a[i + j])
            i += 8
        res = f(f(f(r[0] + r[1]) + f(r[2] + r[3])) + f(f(r[4] + r[5]) + f(r[6] + r[7])))
        while i < n:
            res = f(res + a[i])
            i += 1
        return res
    n2 = n // 2
    n2 -= n2 % 8
    return f(np_pairwise_sum_f32(a[:n2]) + np_pairwise_sum_f32(a[n2:]))


def seq_priority_value(p: np.ndarray, eta: float = 0.9) -> np.float32:
    """eta*max(p) + (1-eta)*mean(p) in float32, as numpy 2.x evaluates
    ``self.eta * priority.max() + (1 - self.eta) * priority.mean()`` (replay_memory.py:188):
    the Python floats ``eta`` and ``1-eta`` are rounded to f32, both products and the sum
    are separate f32 roundings (no FMA), mean = pairwise_sum / n in f32."""
    f = np.float32
    p = np.ascontiguousarray(p, dtype=np.float32)
    n = p.shape[0]
    mx = f(p.max())
    mean = f(np_pairwise_sum_f32(p) / f(n))
    return f(f(f(eta) * mx) + f(f(1 - eta) * mean))


# --------------------------------------------------------------------------------------
# priority updates
# --------------------------------------------------------------------------------------
def update_priority(priority: np.ndarray, index: np.ndarray, p: np.ndarray) -> None:
    """priority[index] = p with numpy fancy-assignment semantics: on duplicate slots the
    LAST occurrence in flattened order wins (replay_memory.py:175-176)."""
    index = np.asarray(index).reshape(-1)
    p = np.asarray(p, dtype=np.float32).reshape(-1)
    for k in range(index.shape[0]):  # explicit sequential order = the contract
        priority[int(index[k])] = p[k]


def update_sequence_priority(priority, sequence_priority, is_seq_start, seq_index,
                             seq_len: int, eta: float = 0.9,
                             update_pre_next_seq_priority: bool = False) -> None:
    """Sequential restatement of replay_memory.py:184-213, quirks included:

    * own slot: NON-wrapping slice priority[idx:idx+T] (truncated at the array end);
    * "previous": nearest earlier start within T slots (wrapping probe), written at
      ``idx-i`` (negative -> numpy wraps) from a wrapping window;
    * "next": probes forward for a start at ``idx+i`` but then uses ``idx-i`` as the slot
      (replay_memory.py:207), again with a wrapping window.
    """
    M = priority.shape[0]
    T = seq_len
    for idx in np.asarray(seq_index).reshape(-1):
        idx = int(idx)
        sequence_priority[idx] = seq_priority_value(priority[idx: idx + T], eta)
        if not update_pre_next_seq_priority:
            continue
        for direction in (-1, +1):
            hit = None
            for i in range(1, T + 1):
                if is_seq_start[(idx + direction * i) % M] == 1:
                    hit = idx - i  # both branches write at idx - i
                    break
            if hit is None:
                continue
            window = (np.arange(hit, hit + T) % M)
            sequence_priority[hit] = seq_priority_value(priority[window], eta)


# --------------------------------------------------------------------------------------
# reference sampling semantics (inverse CDF)
# --------------------------------------------------------------------------------------
def ref_sample_indices(sequence_priority, is_seq_start, u: np.ndarray) -> np.ndarray:
    """Slots the reference's ``sample()`` returns when ``torch.multinomial`` consumes the
    float64 uniforms ``u`` (replay_memory.py:225-232).

    population = slots with is_seq_start==1 in ascending order; weights = their f32
    sequence_priority divided by the f32 (numpy pairwise) sum, widened to f64; multinomial
    with replacement = sequential f64 cumsum, normalised by its total (last bin forced to
    1), first bin whose cdf >= u.
    """
    starts = np.nonzero(np.asarray(is_seq_start) == 1)[0]
    p = np.asarray(sequence_priority, dtype=np.float32)[starts]
    w = (p / np.sum(p)).astype(np.float64)
    cdf = np.cumsum(w)  # sequential double accumulation
    cdf = cdf / cdf[-1]
    cdf[-1] = 1.0
    j = np.searchsorted(cdf, np.asarray(u, dtype=np.float64), side="left")
    return starts[j].astype(np.int64)


# --------------------------------------------------------------------------------------
# the 32-ary f64 sum-tree used by the CUDA path
# --------------------------------------------------------------------------------------
def _hs_scan(v: np.ndarray) -> np.ndarray:
    """Inclusive Hillis-Steele scan over the last axis (length 32), f64, in exactly the
    association the warp-shuffle scan uses: for d in 1,2,4,8,16: v[i] += v[i-d] (i>=d)."""
    v = np.array(v, dtype=np.float64, copy=True)
    for d in (1, 2, 4, 8, 16):
        shifted = np.zeros_like(v)
        shifted[..., d:] = v[..., :-d]
        add = v + shifted
        v[..., d:] = add[..., d:]
    return v


class SumTree:
    """32-ary sum tree over f32 leaves with f64 internal nodes.

    level 0 = leaves (f32, ``sequence_priority`` masked by ``is_seq_start``); level k+1 node
    j = last element of the Hillis-Steele scan of level-k children [32j, 32j+32)
    (zero padded).  Levels are built until one has <= 32 nodes; ``root`` is the scan
    total of that top level.
    """

    def __init__(self, leaves_f32: np.ndarray):
        self.leaves = np.ascontiguousarray(leaves_f32, dtype=np.float32)
        self.levels = []  # internal levels, bottom-up (f64)
        cur = self.leaves.astype(np.float64)
        while True:
            n = cur.shape[0]
            if n <= FAN and self.levels:
                break
            nn = (n + FAN - 1) // FAN
            pad = np.zeros(nn * FAN, dtype=np.float64)
            pad[:n] = cur
            nxt = _hs_scan(pad.reshape(nn, FAN))[:, -1].copy()
            self.levels.append(nxt)
            cur = nxt
            if nn <= FAN:
                break
        top = np.zeros(FAN, dtype=np.float64)
        top[: cur.shape[0]] = cur
        self.root = float(_hs_scan(top)[-1])

    @staticmethod
    def level_sizes(M: int):
        sizes = []
        n = M
        while True:
            n = (n + FAN - 1) // FAN
            sizes.append(n)
            if n <= FAN:
                break
        return sizes

    def _children(self, level: int, node: int) -> np.ndarray:
        """f64 values of the 32 children of `node` at internal level `level`
        (level == -1: the virtual root whose children are the top level)."""
        src = self.leaves if level == 0 else self.levels[level - 1]
        out = np.zeros(FAN, dtype=np.float64)
        lo = node * FAN
        hi = min(lo + FAN, src.shape[0])
        if hi > lo:
            out[: hi - lo] = src[lo:hi]
        return out

    def update_leaf(self, slot: int, value) -> None:
        """Write a leaf and recompute its ancestors (each = scan total of its children)."""
        self.leaves[slot] = np.float32(value)
        node = slot
        for lvl in range(len(self.levels)):
            node //= FAN
            self.levels[lvl][node] = _hs_scan(self._children(lvl, node))[-1]
        top = np.zeros(FAN, dtype=np.float64)
        top[: self.levels[-1].shape[0]] = self.levels[-1]
        self.root = float(_hs_scan(top)[-1])

    @staticmethod
    def _pick(vals: np.ndarray, target: float):
        """Descent rule shared with the kernel: first lane with value>0 and inclusive
        prefix >= target; if none, the last lane with value>0 (lane 0 if all zero).
        Returns (lane, target - exclusive_prefix(lane))."""
        pre = _hs_scan(vals)
        lane = -1
        for j in range(FAN):
            if vals[j] > 0.0 and pre[j] >= target:
                lane = j
                break
        if lane < 0:
            pos = np.nonzero(vals > 0.0)[0]
            lane = int(pos[-1]) if pos.size else 0
        excl = pre[lane - 1] if lane > 0 else 0.0
        return lane, target - excl

    def sample(self, u: np.ndarray) -> np.ndarray:
        out = np.empty(len(u), dtype=np.int64)
        nlev = len(self.levels)
        for k, uk in enumerate(np.asarray(u, dtype=np.float64)):
            target = float(uk) * self.root
            top = np.zeros(FAN, dtype=np.float64)
            top[: self.levels[-1].shape[0]] = self.levels[-1]
            node, target = self._pick(top, target)
            for lvl in range(nlev - 1, -1, -1):
                lane, target = self._pick(self._children(lvl, node), target)
                node = node * FAN + lane
            out[k] = node
        return out


def masked_leaves(sequence_priority, is_seq_start) -> np.ndarray:
    sp = np.asarray(sequence_priority, dtype=np.float32)
    return np.where(np.asarray(is_seq_start) == 1, sp, np.float32(0.0)).astype(np.float32)


# --------------------------------------------------------------------------------------
# slab gather
# --------------------------------------------------------------------------------------
def gather_batch(mem: dict, seq_index: np.ndarray, *, M: int, n_step: int, burn_in: int,
                 learn: int, cell: int) -> dict:
    """What ``ReplayMemory.sample`` assembles after index selection (replay_memory.py:233-262),
    for the only stacking mode the reference uses (stack_count == n_stacks//action_repeat == 1,
    so ``get_stacked_state(i)`` is ``state[i]``, replay_memory.py:215-222).  Frames are
    returned as f32/255 like the reference; layouts are the reference's ([T,B,...])."""
    seq_index = np.asarray(seq_index, dtype=np.int64)
    T = burn_in + learn
    nxt = (seq_index + n_step) % M
    steps = np.arange(T)[:, None]
    idx_state = (seq_index[None, :] + steps) % M
    idx_next = (nxt[None, :] + steps) % M
    lsteps = np.arange(learn)[:, None]
    idx_learn = (seq_index[None, :] + burn_in + lsteps) % M
    b = {}
    b["state"] = mem["state"][idx_state].astype(np.float32) / np.float32(255.0)
    b["next_state"] = mem["state"][idx_next].astype(np.float32) / np.float32(255.0)
    b["hs"] = mem["hs_cs"][seq_index, :cell].copy()
    b["cs"] = mem["hs_cs"][seq_index, cell:].copy()
    b["target_hs"] = mem["target_hs_cs"][nxt, :cell].copy()
    b["target_cs"] = mem["target_hs_cs"][nxt, cell:].copy()
    b["action"] = mem["action"][idx_learn].astype(np.int64)
    b["reward"] = mem["reward"][idx_learn].astype(np.float32)
    b["done"] = mem["done"][idx_learn].astype(np.float32)
    return dict(batch=b, seq_index=seq_index, index=idx_state)


# --------------------------------------------------------------------------------------
# ring buffer
# --------------------------------------------------------------------------------------
KEYS = ("state", "hs_cs", "target_hs_cs", "action", "reward", "done", "stack_count",
        "priority", "sequence_priority", "is_seq_start")


def alloc_memory(M: int, cell: int, frame=(4, 84, 84)) -> dict:
    """Same ten keys / dtypes / shapes as replay_memory.py:76-86."""
    return {
        "state": np.zeros((M, *frame), np.uint8),
        "hs_cs": np.zeros((M, 2 * cell), np.float32),
        "target_hs_cs": np.zeros((M, 2 * cell), np.float32),
        "action": np.zeros((M, 1), np.int8),
        "reward": np.zeros((M, 1), np.float32),
        "done": np.zeros((M, 1), np.float32),
        "stack_count": np.zeros((M,), np.int8),
        "priority": np.zeros((M,), np.float32),
        "sequence_priority": np.zeros((M,), np.float32),
        "is_seq_start": np.zeros((M,), np.uint8),
    }


class ReplayOracle:
    """numpy ring with the reference's write path (replay_memory.py:93-119)."""

    def __init__(self, M, cell=256, n_step=3, burn_in=10, learn=10, eta=0.9):
        self.M, self.cell, self.n_step = M, cell, n_step
        self.burn_in, self.learn, self.eta = burn_in, learn, eta
        self.index = 0
        self.memory = alloc_memory(M, cell)

    @property
    def T(self):
        return self.burn_in + self.learn

    @property
    def size(self):
        return min(self.index, self.M)

    def add(self, state, hs, cs, target_hs, target_cs, action, reward, done, stack_count,
            priority):
        i = self.index % self.M
        m = self.memory
        m["state"][i] = state * 255  # float -> u8 truncation, as replay_memory.py:95
        m["hs_cs"][i, : self.cell] = hs
        m["hs_cs"][i, self.cell:] = cs
        m["target_hs_cs"][i, : self.cell] = target_hs
        m["target_hs_cs"][i, self.cell:] = target_cs
        m["action"][i] = action
        m["reward"][i] = reward
        m["done"][i] = 1 if done else 0
        m["stack_count"][i] = stack_count
        m["priority"][i] = priority
        self.index += 1

    def extend(self, dump: dict):
        rows = dump["state"].shape[0]
        slots = (self.index % self.M + np.arange(rows)) % self.M
        for k in KEYS:
            self.memory[k][slots] = dump[k]
        self.index += rows

    def update_priority(self, index, p):
        update_priority(self.memory["priority"], index, p)

    def update_sequence_priority(self, index, flag=False):
        update_sequence_priority(self.memory["priority"], self.memory["sequence_priority"],
                                 self.memory["is_seq_start"], index, self.T, self.eta, flag)

    def sample_indices(self, u):
        return ref_sample_indices(self.memory["sequence_priority"],
                                  self.memory["is_seq_start"], u)

    def gather(self, seq_index):
        return gather_batch(self.memory, seq_index, M=self.M, n_step=self.n_step,
                            burn_in=self.burn_in, learn=self.learn, cell=self.cell)


def synth_fill(rep: ReplayOracle, seed=1, fill_frac=0.8, stride=None, done_period=997):
    """SYN-R2D2 of SURVEY.md section 8(d): deterministic synthetic replay contents."""
    rng = np.random.RandomState(seed)
    M, H = rep.M, rep.cell
    N = int(fill_frac * M)
    m = rep.memory
    m["state"][:N] = rng.randint(0, 256, (N, 4, 84, 84)).astype(np.uint8)
    m["hs_cs"][:N] = (rng.randn(N, 2 * H) * 0.1).astype(np.float32)
    m["target_hs_cs"][:N] = (rng.randn(N, 2 * H) * 0.1).astype(np.float32)
    m["action"][:N, 0] = rng.randint(6, size=N).astype(np.int8)
    m["reward"][:N, 0] = rng.randn(N).astype(np.float32)
    m["done"][:N, 0] = (np.arange(N) % done_period == done_period - 1).astype(np.float32)
    m["stack_count"][:N] = 1
    m["priority"][:N] = (rng.rand(N) + 0.01).astype(np.float32)
    rep.index = N
    stride = stride or rep.learn
    starts = np.arange(0, max(N - rep.T - rep.n_step, 1), stride)
    rep.update_sequence_priority(starts)
    m["is_seq_start"][starts] = 1
    return starts
