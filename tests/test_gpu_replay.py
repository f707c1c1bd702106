"""GPU parity: replay kernels (through the C ABI) against the oracle and the golden vectors.
Bit-exact for indices, priorities and bytes."""
import os

import numpy as np
import pytest
import torch

from oracle import replay_oracle as RO
from test_oracle_golden import sampling_inputs
from util import gpu_mem_np, make_oracle_replay, mirror_to_gpu

pytestmark = pytest.mark.gpu


def tree_on_gpu(leaves, mask, dev):
    from pytorch_r2d2_b200 import _lib
    from pytorch_r2d2_b200.replay_memory import tree_layout
    M = leaves.shape[0]
    sizes, offs, total = tree_layout(M)
    sp = torch.from_numpy(leaves).to(dev)
    st = torch.from_numpy(mask).to(dev)
    nodes = torch.zeros(total, dtype=torch.float64, device=dev)
    _lib.call("r2d2_tree_build", _lib.ptr(sp, "f32"), _lib.ptr(st, "u8"), M, _lib.ptr(nodes, "f64"),
              _lib.stream_ptr())
    return sp, st, nodes, sizes, offs


def tree_sample_gpu(sp, st, nodes, u):
    from pytorch_r2d2_b200 import _lib
    ud = torch.from_numpy(np.asarray(u, np.float64)).to(sp.device)
    out = torch.empty(ud.numel(), dtype=torch.int64, device=sp.device)
    _lib.call("r2d2_tree_sample", _lib.ptr(sp, "f32"), _lib.ptr(st, "u8"), sp.numel(),
              _lib.ptr(nodes, "f64"), _lib.ptr(ud, "f64"), ud.numel(), _lib.ptr(out, "i64"),
              _lib.stream_ptr())
    return out.cpu().numpy()


@pytest.fixture(scope="module")
def gold(golden_dir):
    return {n: np.load(os.path.join(golden_dir, n + ".npz")) for n in ("sampling", "priority")}


@pytest.mark.parametrize("name", ["n37", "n1000", "n100000", "x1024", "x65536"])
def test_tree_build_and_sample_vs_oracle_tree(cuda_device, gold, name):
    g = gold["sampling"]
    leaves, mask = sampling_inputs(g, name)
    sp, st, nodes, sizes, offs = tree_on_gpu(leaves, mask, cuda_device)
    t = RO.SumTree(RO.masked_leaves(leaves, mask))
    nd = nodes.cpu().numpy()
    for lvl, (sz, off) in enumerate(zip(sizes, offs)):       # tier (i): node sums bit-exact
        assert np.array_equal(nd[off:off + sz].view(np.uint64), t.levels[lvl].view(np.uint64)), lvl
    u = g[f"{name}/u"]
    got = tree_sample_gpu(sp, st, nodes, u)
    assert np.array_equal(got[:512], t.sample(u[:512]))      # tier (i): descent bit-exact
    ref = g[f"{name}/seq_index"]                              # tiers (ii)/(iii): the reference's draws
    assert int((got == ref).sum()) == len(ref), f"{(got != ref).sum()} draws differ from the reference"


@pytest.mark.parametrize("name", ["x4m", "r4m"])
def test_tree_4m_leaves_vs_reference(cuda_device, gold, name):
    """BASELINE config 3 size: 4M sequence starts; GPU draws == the reference's sample()."""
    g = gold["sampling"]
    leaves, mask = sampling_inputs(g, name)
    sp, st, nodes, _, _ = tree_on_gpu(leaves, mask, cuda_device)
    got = tree_sample_gpu(sp, st, nodes, g[f"{name}/u"])
    ref = g[f"{name}/seq_index"]
    assert int((got == ref).sum()) == len(ref)
    # size-independent properties: only starts with positive weight are drawn; root == f64 sum
    assert np.all(mask[got] == 1) and np.all(leaves[got] > 0)
    sizes, offs, _ = __import__("pytorch_r2d2_b200.replay_memory", fromlist=["x"]).tree_layout(len(leaves))
    top = nodes[offs[-1]: offs[-1] + sizes[-1]].sum().item()
    assert abs(top - leaves.astype(np.float64).sum()) <= 1e-9 * top


def test_tree_edge_cases(cuda_device):
    rng = np.random.RandomState(3)
    for M in (1, 5, 31, 32, 33, 1023, 1025, 40000):
        leaves = rng.rand(M).astype(np.float32)
        mask = (rng.rand(M) < 0.3).astype(np.uint8)
        mask[rng.randint(M)] = 1
        sp, st, nodes, _, _ = tree_on_gpu(leaves, mask, cuda_device)
        u = np.concatenate([rng.rand(200), [0.0, 1.0 - 2.0 ** -53, 0.5]])
        got = tree_sample_gpu(sp, st, nodes, u)
        want = RO.SumTree(RO.masked_leaves(leaves, mask)).sample(u)
        assert np.array_equal(got, want), M
        assert np.all(mask[got] == 1), M


def test_tree_update_equals_rebuild(cuda_device):
    from pytorch_r2d2_b200 import _lib
    rng = np.random.RandomState(4)
    M = 70000
    leaves = rng.rand(M).astype(np.float32)
    mask = (rng.rand(M) < 0.5).astype(np.uint8)
    sp, st, nodes, _, _ = tree_on_gpu(leaves, mask, cuda_device)
    slots = rng.randint(0, M, 300).astype(np.int64)
    slots[::7] = -1                                         # skipped entries
    valid = slots[slots >= 0]
    sp[torch.from_numpy(valid).to(sp.device)] = torch.from_numpy(rng.rand(len(valid)).astype(np.float32)).to(sp.device)
    sl = torch.from_numpy(slots).to(sp.device)
    _lib.call("r2d2_tree_update", _lib.ptr(sp, "f32"), _lib.ptr(st, "u8"), M, _lib.ptr(nodes, "f64"),
              _lib.ptr(sl, "i64"), len(slots), _lib.stream_ptr())
    _, _, nodes2, _, _ = tree_on_gpu(sp.cpu().numpy(), mask, cuda_device)
    assert torch.equal(nodes, nodes2)


@pytest.mark.parametrize("use_tma", [False, True])
def test_gather_vs_oracle(cuda_device, use_tma):
    M, H, NS, BI, LN, B = 96, 256, 2, 3, 4, 6
    rep, starts = make_oracle_replay(M, H, NS, BI, LN, seed=31)
    g = mirror_to_gpu(rep, B)
    g.use_tma_gather = use_tma
    # includes sequences that wrap around the ring end and a duplicate
    seq = np.array([0, 90, 93, 95, 45, 45], dtype=np.int64)
    ws = g.sample_slab(seq_index=seq)
    torch.cuda.synchronize()
    want = rep.gather(seq)
    S = BI + LN + NS
    slab = ws["slab"].cpu().numpy()
    for b in range(B):
        rows = (seq[b] + np.arange(S)) % M
        assert np.array_equal(slab[b], rep.memory["state"][rows]), b
    wb = want["batch"]
    for k in ("hs", "cs", "target_hs", "target_cs"):
        assert np.array_equal(ws[k].cpu().numpy(), wb[k]), k
    assert np.array_equal(ws["action"].cpu().numpy(), wb["action"][..., 0])
    assert np.array_equal(ws["reward"].cpu().numpy(), wb["reward"][..., 0])
    assert np.array_equal(ws["done"].cpu().numpy(), wb["done"][..., 0])
    assert np.array_equal(ws["index"].cpu().numpy(), want["index"])


def test_sample_api_matches_reference_semantics(cuda_device):
    """ReplayMemory.sample(): same uniforms as torch.multinomial -> same indices and batch as the
    oracle (itself pinned to the reference in make_golden.py)."""
    M, H, NS, BI, LN, B = 64, 256, 2, 3, 4, 4
    rep, _ = make_oracle_replay(M, H, NS, BI, LN, seed=21)
    g = mirror_to_gpu(rep, B)
    torch.manual_seed(1000)
    batch, seq_index, index = g.sample()
    torch.manual_seed(1000)
    u = torch.rand(B, dtype=torch.float64).numpy()
    want_idx = rep.sample_indices(u)
    assert np.array_equal(seq_index, want_idx)
    want = rep.gather(want_idx)
    assert np.array_equal(index, want["index"])
    for k, v in want["batch"].items():
        got = batch[k].cpu().numpy()
        assert got.shape == v.shape, (k, got.shape, v.shape)
        assert np.array_equal(got, v.astype(got.dtype)), k
    assert batch["action"].dtype == torch.int64 and batch["state"].dtype == torch.float32


def test_ingest_extend_wraps(cuda_device):
    from pytorch_r2d2_b200 import ReplayMemory
    M, H = 50, 256
    rep = RO.ReplayOracle(M, H)
    g = ReplayMemory(M, 4, 3, cell_size=H)
    rng = np.random.RandomState(9)
    for rows in (20, 25, 17, 1):                              # third dump wraps the ring
        dump = RO.alloc_memory(rows, H)
        dump["state"][:] = rng.randint(0, 256, dump["state"].shape)
        for k in ("hs_cs", "target_hs_cs", "reward", "priority", "sequence_priority"):
            dump[k][:] = rng.randn(*dump[k].shape)
        dump["action"][:] = rng.randint(0, 6, dump["action"].shape)
        dump["done"][:] = rng.randint(0, 2, dump["done"].shape)
        dump["stack_count"][:] = 1
        dump["is_seq_start"][:] = rng.randint(0, 2, rows)
        rep.extend(dump)
        g.extend(dump)
    assert g.index == rep.index and g.size == rep.size
    for k in RO.KEYS:
        assert np.array_equal(gpu_mem_np(g, k), rep.memory[k]), k


def test_add_matches_reference_semantics(cuda_device):
    from pytorch_r2d2_b200 import ReplayMemory
    M, H = 8, 256
    rep = RO.ReplayOracle(M, H)
    g = ReplayMemory(M, 4, 3, cell_size=H)
    rng = np.random.RandomState(2)
    for i in range(11):                                        # wraps
        st = rng.randint(0, 256, (4, 84, 84)).astype(np.float32) / np.float32(255.)
        args = (st, rng.randn(H).astype(np.float32), rng.randn(H).astype(np.float32),
                rng.randn(H).astype(np.float32), rng.randn(H).astype(np.float32),
                int(rng.randint(6)), float(rng.randn()), bool(i % 3 == 0), 1, float(rng.rand()))
        rep.add(*args)
        g.add(*args)
    for k in RO.KEYS:
        assert np.array_equal(gpu_mem_np(g, k), rep.memory[k]), k


@pytest.mark.parametrize("name", ["regular", "irregular", "near_end", "noflag", "dense"])
def test_priority_updates_vs_golden(cuda_device, gold, name):
    """update_priority + update_sequence_priority == the reference's arrays, bit for bit."""
    from pytorch_r2d2_b200 import ReplayMemory
    g = gold["priority"]
    M, T = int(g[f"{name}/M"]), int(g[f"{name}/T"])
    r = ReplayMemory(M, 4, 1, cell_size=32)
    r.burn_in_length, r.learning_length, r.sequence_length = T // 2, T - T // 2, T
    raw = lambda k: dict.__getitem__(r.memory, k)  # noqa: E731
    raw("priority").copy_(torch.from_numpy(g[f"{name}/priority0"]))
    raw("sequence_priority").copy_(torch.from_numpy(g[f"{name}/seqprio0"]))
    r.memory["is_seq_start"][torch.from_numpy(g[f"{name}/starts"]).cuda()] = 1   # direct poke, like actor.py:167
    r.update_priority(g[f"{name}/scatter_idx"], g[f"{name}/scatter_val"])
    assert np.array_equal(gpu_mem_np(r, "priority"), g[f"{name}/priority1"])
    r._ensure_tree()
    r.update_sequence_priority(g[f"{name}/seq_index"], bool(g[f"{name}/flag"]))
    got = gpu_mem_np(r, "sequence_priority")
    assert np.array_equal(got.view(np.uint32), g[f"{name}/seqprio1"].view(np.uint32))
    # incremental tree == rebuilt tree
    nodes = r._nodes.clone()
    r.rebuild_tree()
    assert torch.equal(nodes, r._nodes)


def test_learner_priority_scatter_duplicates(cuda_device):
    """r2d2_update_priority_seq == numpy fancy assignment on overlapping / duplicate sequences."""
    from pytorch_r2d2_b200 import ReplayMemory
    M, BI, LN, B = 40, 3, 5, 7
    r = ReplayMemory(M, B, 1, cell_size=32)
    r.burn_in_length, r.learning_length, r.sequence_length = BI, LN, BI + LN
    rng = np.random.RandomState(6)
    pr0 = rng.rand(M).astype(np.float32)
    dict.__getitem__(r.memory, "priority").copy_(torch.from_numpy(pr0))
    seq = np.array([0, 2, 2, 35, 38, 4, 33], dtype=np.int64)    # overlaps, duplicate, ring wrap
    p = rng.rand(LN, B).astype(np.float32)
    index = np.stack([(seq + s) % M for s in range(BI + LN)])
    want = pr0.copy()
    want[index[BI:].reshape(-1)] = p.reshape(-1)                 # learner.py:103
    r.update_priority_learn_window(torch.from_numpy(seq).cuda(), torch.from_numpy(p).cuda())
    assert np.array_equal(gpu_mem_np(r, "priority"), want)


def test_full_size_gather_roundtrip(cuda_device):
    """BASELINE size (B=64, T=80, n=5): every slab row equals its ring row (checksum of sums)."""
    from pytorch_r2d2_b200 import ReplayMemory
    M, B, T, n = 4000, 64, 80, 5
    r = ReplayMemory(M, B, n, cell_size=512)
    r.burn_in_length, r.learning_length, r.sequence_length = 40, 40, 80
    st = dict.__getitem__(r.memory, "state")
    st.random_(0, 256)
    seq = torch.randint(0, M, (B,), device="cuda")
    seq[0] = M - 3                                               # wraps
    for tma in (False, True):
        r.use_tma_gather = tma
        ws = r.sample_slab(seq_index=seq)
        rows = (seq[:, None] + torch.arange(T + n, device="cuda")[None]) % M
        assert torch.equal(ws["slab"], st[rows]), tma


def test_overlapped_extend_is_visible_to_next_draw(cuda_device):
    """extend(overlap=True) runs on the ingest stream; the following draw/gather must see it."""
    from pytorch_r2d2_b200 import ReplayMemory
    M, H = 64, 256
    rep = RO.ReplayOracle(M, H, 2, 3, 4)
    g = ReplayMemory(M, 4, 2, cell_size=H)
    g.burn_in_length, g.learning_length, g.sequence_length = 3, 4, 7
    rng = np.random.RandomState(3)
    for rows in (30, 30, 20):
        dump = RO.alloc_memory(rows, H)
        dump["state"][:] = rng.randint(0, 256, dump["state"].shape)
        for k in ("hs_cs", "target_hs_cs", "reward", "priority"):
            dump[k][:] = rng.randn(*dump[k].shape)
        dump["sequence_priority"][:] = rng.rand(rows) + 0.1
        dump["is_seq_start"][:] = (np.arange(rows) % 5 == 0)
        rep.extend(dump)
        pinned = {k: torch.from_numpy(v).pin_memory() for k, v in dump.items()}
        g.extend(pinned, overlap=True)
        u = rng.rand(4)
        ws = g.sample_slab(u=torch.from_numpy(u))
        si = rep.sample_indices(u)
        assert np.array_equal(ws["seq_index"].cpu().numpy(), si)
        rows_idx = (si[:, None] + np.arange(9)[None]) % M
        assert np.array_equal(ws["slab"].cpu().numpy(), rep.memory["state"][rows_idx])


@pytest.mark.parametrize("B", [300, 1024, 4096])
def test_large_batch_priority_path_vs_oracle(cuda_device, B):
    """BASELINE config 3 batch sizes take the grid-wide kernels (O(n) claim/commit scatter, one
    warp per sequence refresh, per-level tree update): same arrays as the oracle, bit for bit,
    with duplicate and overlapping sequences, starts near the ring end, and a generic
    update_priority call large enough for the claim/commit form."""
    from pytorch_r2d2_b200 import ReplayMemory
    M, BI, LN = 20000, 6, 14
    T = BI + LN
    rng = np.random.RandomState(B)
    rep = RO.ReplayOracle(M, 32, 1, BI, LN)
    rep.memory["priority"][:] = rng.rand(M).astype(np.float32) + 0.01
    starts = np.unique(np.concatenate([np.arange(0, M, 7), [M - 1, M - 5, M - T, M - T + 1]]))
    rep.memory["is_seq_start"][starts] = 1
    rep.memory["sequence_priority"][:] = rng.rand(M).astype(np.float32)
    rep.index = M
    g = ReplayMemory(M, 4, 1, cell_size=32)
    g.burn_in_length, g.learning_length, g.sequence_length = BI, LN, T
    g.extend({k: v.copy() for k, v in rep.memory.items()})
    g._ensure_tree()
    seq = rng.choice(starts, B).astype(np.int64)                  # duplicates and overlaps
    seq[:4] = [M - 1, M - 5, M - T + 1, M - 1]
    p = rng.rand(LN, B).astype(np.float32)
    index = np.stack([(seq + s) % M for s in range(T)])
    rep.update_priority(index[BI:].reshape(-1), p.reshape(-1))
    rep.update_sequence_priority(seq, True)
    g.update_priority_learn_window(torch.from_numpy(seq).cuda(), torch.from_numpy(p).cuda())
    g.update_sequence_priority(torch.from_numpy(seq).cuda(), True)
    assert np.array_equal(gpu_mem_np(g, "priority").view(np.uint32), rep.memory["priority"].view(np.uint32))
    assert np.array_equal(gpu_mem_np(g, "sequence_priority").view(np.uint32),
                          rep.memory["sequence_priority"].view(np.uint32))
    nodes = g._nodes.clone()
    g.rebuild_tree()
    assert torch.equal(nodes, g._nodes)
    # generic scatter, n > 2048 -> claim/commit; negative (numpy-style) indices included
    idx = rng.randint(-M, M, 5000).astype(np.int64)
    val = rng.rand(5000).astype(np.float32)
    rep.update_priority(idx, val)
    g.update_priority(idx, val)
    assert np.array_equal(gpu_mem_np(g, "priority").view(np.uint32), rep.memory["priority"].view(np.uint32))
