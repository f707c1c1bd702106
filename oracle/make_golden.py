"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):

    python oracle/make_golden.py

The reference is imported as-is; the single missing third-party import (`fasteners`, used
only by the file hand-off `save/load`, replay_memory.py:127,157) is stubbed with an empty
module.  Nothing from the reference is copied into the repo: only its OUTPUTS on seeded
synthetic inputs are stored (small .npz files), together with this script.
Synthetic inputs are regenerated inside the tests from the same seeds.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)

from oracle import replay_oracle as RO  # noqa: E402
from oracle import net_oracle as NO  # noqa: E402


def import_reference(path="/root/reference"):
    if "fasteners" not in sys.modules:
        sys.modules["fasteners"] = types.ModuleType("fasteners")
    if path not in sys.path:
        sys.path.insert(0, path)
    import replay_memory as ref_replay  # noqa
    import model as ref_model  # noqa
    import learner as ref_learner  # noqa
    return ref_replay, ref_model, ref_learner


# ------------------------------------------------------------------ sampling goldens
def tiny_ref_replay(ref_replay, M, B, n_step, bi, ln):
    """Reference ReplayMemory with 1x1 frames and 1-wide cells so that the literal
    sample() runs at millions of slots (index selection is independent of both)."""
    r = ref_replay.ReplayMemory(M, B, n_step, state_size=(1, 1), cell_size=1)
    r.burn_in_length, r.learning_length, r.sequence_length = bi, ln, bi + ln
    return r


def sampling_case(ref_replay, name, N, B, seed_prio, seed_u, exact, stride=1):
    """N sequence starts (every `stride` slots)."""
    M = N * stride
    r = tiny_ref_replay(ref_replay, M, B, 1, 1, 1)
    rng = np.random.RandomState(seed_prio)
    if exact:
        # integer priorities in [1,3], total nudged to exactly 2N with N a power of two:
        # every f32/f64 operation in the reference's path is then exact.
        p = rng.randint(1, 4, N).astype(np.int64)
        diff = int(2 * N - p.sum())
        i = 0
        while diff != 0:
            step = 1 if diff > 0 else -1
            if 1 <= p[i] + step <= 3:
                p[i] += step
                diff -= step
            i = (i + 1) % N
        assert p.sum() == 2 * N
        p = p.astype(np.float32)
    else:
        p = ((rng.rand(N).astype(np.float32) + np.float32(1e-3)) ** np.float32(0.6)).astype(np.float32)
    starts = np.arange(N) * stride
    r.memory["sequence_priority"][starts] = p
    r.memory["is_seq_start"][starts] = 1
    r.index = M
    torch.manual_seed(seed_u)
    _, seq_index, _ = r.sample("cpu")
    # the uniforms multinomial consumed (same generator state):
    torch.manual_seed(seed_u)
    u = torch.rand(B, dtype=torch.float64).numpy()
    return dict(name=name, N=N, B=B, stride=stride, seed_prio=seed_prio, seed_u=seed_u,
                exact=int(exact), seq_index=np.asarray(seq_index, dtype=np.int64), u=u), p


def make_sampling(ref_replay):
    out = {}
    cases = [("n37", 37, 64, 11, 101, False, 3), ("n1000", 1000, 256, 12, 102, False, 2),
             ("n100000", 100000, 1024, 13, 103, False, 1),
             ("x1024", 1 << 10, 64, 7, 201, True, 1), ("x65536", 1 << 16, 4096, 7, 202, True, 1),
             ("x4m", 1 << 22, 4096, 7, 203, True, 1), ("r4m", 4000000, 4096, 0, 123, False, 1)]
    for name, N, B, sp, su, exact, stride in cases:
        c, p = sampling_case(ref_replay, name, N, B, sp, su, exact, stride)
        # oracle restatement of the reference CDF must agree with the literal reference
        leaves = np.zeros(N * stride, np.float32)
        leaves[np.arange(N) * stride] = p
        mask = np.zeros(N * stride, np.uint8)
        mask[np.arange(N) * stride] = 1
        mine = RO.ref_sample_indices(leaves, mask, c["u"])
        agree = int((mine == c["seq_index"]).sum())
        tree = RO.SumTree(RO.masked_leaves(leaves, mask)).sample(c["u"])
        agree_tree = int((tree == c["seq_index"]).sum())
        print(f"[sampling] {name}: restated-CDF == reference {agree}/{B}; 32-ary f64 tree == reference {agree_tree}/{B}")
        assert agree == B, "u-extraction or CDF restatement is wrong"
        if exact:
            assert agree_tree == B
        c["tree_agree"] = agree_tree
        for k, v in c.items():
            if k != "name":
                out[f"{name}/{k}"] = np.asarray(v)
    np.savez_compressed(os.path.join(GOLD, "sampling.npz"), **out)


# ------------------------------------------------------------------ priority goldens
def make_priority(ref_replay):
    out = {}
    cases = {
        # name: (M, T, starts, seq_index, flag)
        "regular": (200, 20, list(range(0, 160, 10)), [30, 50, 50, 0, 150], True),
        "irregular": (200, 20, [0, 7, 19, 33, 34, 60, 95, 96, 140, 171], [33, 60, 7, 171, 0, 96], True),
        "near_end": (120, 20, [0, 10, 90, 100, 105, 110, 115, 119], [110, 119, 105, 0, 115], True),
        "noflag": (120, 20, [0, 10, 90, 100, 110], [110, 0, 100], False),
        "dense": (64, 7, list(range(0, 60, 3)), [57, 3, 0, 3, 60 - 3], True),
    }
    for name, (M, T, starts, seq_index, flag) in cases.items():
        r = tiny_ref_replay(ref_replay, M, 4, 1, T // 2, T - T // 2)
        rng = np.random.RandomState(hash(name) % 1000 if False else len(name) * 17 + M)
        pr = (rng.rand(M).astype(np.float32) + np.float32(0.01))
        r.memory["priority"][:] = pr
        r.memory["is_seq_start"][starts] = 1
        sp0 = rng.rand(M).astype(np.float32)
        r.memory["sequence_priority"][:] = sp0
        # scatter with duplicates first (update_priority), then the sequence refresh
        idx = rng.randint(0, M, 3 * T).astype(np.int64)
        val = rng.rand(3 * T).astype(np.float32)
        r.update_priority(idx, val)
        after_scatter = r.memory["priority"].copy()
        r.update_sequence_priority(np.array(seq_index), flag)
        # oracle check
        pr2 = pr.copy()
        RO.update_priority(pr2, idx, val)
        assert np.array_equal(pr2, after_scatter), name
        sp2 = sp0.copy()
        mask = np.zeros(M, np.uint8)
        mask[starts] = 1
        RO.update_sequence_priority(pr2, sp2, mask, np.array(seq_index), T, 0.9, flag)
        assert np.array_equal(sp2.view(np.uint32), r.memory["sequence_priority"].view(np.uint32)), name
        print(f"[priority] {name}: oracle bit-exact vs reference")
        out.update({f"{name}/M": M, f"{name}/T": T, f"{name}/starts": np.array(starts),
                    f"{name}/seq_index": np.array(seq_index), f"{name}/flag": int(flag),
                    f"{name}/priority0": pr, f"{name}/seqprio0": sp0, f"{name}/scatter_idx": idx,
                    f"{name}/scatter_val": val, f"{name}/priority1": after_scatter,
                    f"{name}/seqprio1": r.memory["sequence_priority"].copy()})
    # pairwise-sum restatement vs numpy on many lengths
    rng = np.random.RandomState(5)
    for n in list(range(1, 40)) + [80, 85, 127, 128, 129, 200, 1000]:
        for _ in range(20):
            a = rng.rand(n).astype(np.float32)
            assert RO.np_pairwise_sum_f32(a).view(np.uint32) == np.sum(a).view(np.uint32), n
    print("[priority] pairwise-sum restatement == np.sum on all probed lengths")
    np.savez_compressed(os.path.join(GOLD, "priority.npz"), **out)


# ------------------------------------------------------------------ network goldens
def fill_small(ref_or_oracle_mem, M, H, seed):
    rng = np.random.RandomState(seed)
    m = ref_or_oracle_mem
    N = M
    m["state"][:N] = rng.randint(0, 256, (N, 4, 84, 84)).astype(np.uint8)
    m["hs_cs"][:N] = (rng.randn(N, 2 * H) * 0.1).astype(np.float32)
    m["target_hs_cs"][:N] = (rng.randn(N, 2 * H) * 0.1).astype(np.float32)
    m["action"][:N, 0] = rng.randint(6, size=N).astype(np.int8)
    m["reward"][:N, 0] = rng.randn(N).astype(np.float32)
    m["done"][:N, 0] = (np.arange(N) % 13 == 12).astype(np.float32)
    m["stack_count"][:N] = 1
    m["priority"][:N] = (rng.rand(N) + 0.01).astype(np.float32)


def sub(t, n=2048):
    """Deterministic strided subsample of a tensor (fixtures stay small)."""
    f = t.detach().reshape(-1)
    step = max(1, f.numel() // n)
    return f[::step][:n].numpy().copy()


def make_network(ref_replay, ref_model, ref_learner):
    out = {}
    # ---- init parity: oracle init == reference QNet init for the same seed
    torch.manual_seed(0)
    net = ref_model.QNet("cpu")
    sd = net.state_dict()
    p = NO.init_params(0, 256)
    for k in NO.PARAM_KEYS:
        assert torch.equal(sd[k], p[k]), k
    print("[network] init_params(0) == reference QNet() under manual_seed(0)")

    # ---- forward goldens
    rng = np.random.RandomState(3)
    x5 = torch.from_numpy(rng.randint(0, 256, (3, 2, 4, 84, 84)).astype(np.float32) / np.float32(255))
    x4 = x5[0].clone()
    h0 = torch.from_numpy((rng.randn(2, 256) * 0.1).astype(np.float32))
    c0 = torch.from_numpy((rng.randn(2, 256) * 0.1).astype(np.float32))
    with torch.no_grad():
        net.reset()
        q5, hs5, cs5 = net(x5, True)                   # zero initial state, 5-D
        net.set_state(h0.clone(), c0.clone())
        q4, hs4, cs4 = net(x4, True)                   # given state, 4-D
        q4b, hs4b, cs4b = net(x4, True)                # state carries over
    o5 = NO.qnet_forward(p, x5, None, None)
    o4 = NO.qnet_forward(p, x4, h0, c0)
    o4b = NO.qnet_forward(p, x4, *o4[3])
    for a, b in ((q5, o5[0]), (torch.from_numpy(hs5), o5[1]), (q4, o4[0]), (q4b, o4b[0]),
                 (torch.from_numpy(cs4b), o4b[2])):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-6), (a - b).abs().max()
    print("[network] oracle forward == reference forward (5-D zero state, 4-D set_state, carry-over)")
    out.update({"fwd/seed_x": 3, "fwd/q5": q5.numpy(), "fwd/hs5": hs5, "fwd/cs5": cs5,
                "fwd/q4": q4.numpy(), "fwd/hs4": hs4, "fwd/cs4": cs4, "fwd/q4b": q4b.numpy(),
                "fwd/hs4b": hs4b, "fwd/cs4b": cs4b})

    # ---- one (and three) full reference train() steps on a small config
    M, B, BI, LN, NS, H = 64, 4, 3, 4, 2, 256
    torch.manual_seed(0)
    L = ref_learner.Learner(1, {}, device="cpu")
    L.replay_memory = ref_replay.ReplayMemory(M, B, NS)
    for obj in (L, L.replay_memory):
        obj.burn_in_length, obj.learning_length, obj.sequence_length = BI, LN, BI + LN
    L.bootstrap_steps = NS
    L.batch_size = B
    fill_small(L.replay_memory.memory, M, H, seed=21)
    L.replay_memory.index = M
    starts = np.arange(0, M - (BI + LN) - NS, 3)
    L.replay_memory.update_sequence_priority(starts)
    L.replay_memory.memory["is_seq_start"][starts] = 1
    p0 = {k: v.detach().clone() for k, v in L.net.state_dict().items()}
    for k in NO.PARAM_KEYS:
        assert torch.equal(p0[k], p[k])   # Learner built its net first under seed 0
    pt0 = {k: v.detach().clone() for k, v in L.target_net.state_dict().items()}

    # oracle twin of the replay
    rep = RO.ReplayOracle(M, H, NS, BI, LN)
    fill_small(rep.memory, M, H, seed=21)
    rep.index = M
    rep.update_sequence_priority(starts)
    rep.memory["is_seq_start"][starts] = 1
    assert np.array_equal(rep.memory["sequence_priority"], L.replay_memory.memory["sequence_priority"])

    rec = {"net": [], "tgt": [], "sample": []}
    for tag, mod in (("net", L.net), ("tgt", L.target_net)):
        orig = mod.forward

        def wrapped(*a, _o=orig, _t=tag, **kw):
            r = _o(*a, **kw)
            rec[_t].append((r[0] if isinstance(r, tuple) else r).detach().clone())
            return r
        mod.forward = wrapped
    orig_sample = L.replay_memory.sample

    def wrapped_sample(*a, **kw):
        r = orig_sample(*a, **kw)
        rec["sample"].append(r)
        return r
    L.replay_memory.sample = wrapped_sample

    p_or = {k: v.clone() for k, v in p0.items()}
    pt_or = {k: v.clone() for k, v in pt0.items()}
    sq = {k: torch.zeros_like(v) for k, v in p_or.items()}
    ga = {k: torch.zeros_like(v) for k, v in p_or.items()}
    n_steps = 3
    for step in range(n_steps):
        for k in rec:
            rec[k].clear()
        torch.manual_seed(1000 + step)
        L.train()
        batch, seq_index, index = rec["sample"][0]
        # --- oracle, same uniforms
        torch.manual_seed(1000 + step)
        u = torch.rand(B, dtype=torch.float64).numpy()
        si = rep.sample_indices(u)
        assert np.array_equal(si, seq_index), (si, seq_index)
        g = rep.gather(si)
        for k in g["batch"]:
            assert np.array_equal(np.asarray(batch[k].numpy(), dtype=np.float32).reshape(g["batch"][k].shape),
                                  g["batch"][k].astype(np.float32)), k
        assert np.array_equal(g["index"], index)
        res = NO.train_step(p_or, pt_or, NO.to_torch_batch(g["batch"]), burn_in=BI, n_step=NS)
        q_all_ref, q_next_online_ref, q_next_target_ref = rec["net"][1], rec["net"][2], rec["tgt"][1]
        assert torch.allclose(res["q_all"], q_all_ref, rtol=1e-5, atol=1e-6)
        assert torch.allclose(res["q_next_online"], q_next_online_ref, rtol=1e-5, atol=1e-6)
        assert torch.allclose(res["q_next_target"], q_next_target_ref, rtol=1e-5, atol=1e-6)
        ref_grads = {k: dict(L.net.named_parameters())[k].grad.detach().clone() for k in NO.PARAM_KEYS}
        for k in NO.PARAM_KEYS:
            den = ref_grads[k].abs().max().clamp_min(1e-12)
            err = ((res["grads"][k] - ref_grads[k]).abs().max() / den).item()
            assert err < 1e-5, (k, err)
        # optimiser restatement
        for k in NO.PARAM_KEYS:
            NO.rmsprop_centered(p_or[k], res["grads"][k], sq[k], ga[k])
        ref_params = {k: v.detach().clone() for k, v in L.net.state_dict().items()}
        for k in NO.PARAM_KEYS:
            den = ref_params[k].abs().max()
            assert ((p_or[k] - ref_params[k]).abs().max() / den).item() < 1e-5, k
        # priorities
        rep.update_priority(g["index"][BI:].reshape(-1), res["priority"])
        rep.update_sequence_priority(si, True)
        ref_prio = L.replay_memory.memory["priority"]
        assert np.allclose(rep.memory["priority"], ref_prio, rtol=1e-4, atol=1e-7)
        assert np.allclose(rep.memory["sequence_priority"], L.replay_memory.memory["sequence_priority"],
                           rtol=1e-4, atol=1e-7)
        print(f"[network] train step {step}: oracle == reference (indices exact; q/grads/params/priorities in tolerance)")
        if step == 0:
            out.update({"train/seq_index": np.asarray(seq_index), "train/u": u,
                        "train/q_all": q_all_ref.numpy(), "train/q_next_online": q_next_online_ref.numpy(),
                        "train/q_next_target": q_next_target_ref.numpy(),
                        "train/priority_after": ref_prio.copy(),
                        "train/seqprio_after": L.replay_memory.memory["sequence_priority"].copy()})
            for k in NO.PARAM_KEYS:
                out[f"train/grad_sub/{k}"] = sub(ref_grads[k])
                out[f"train/grad_absmax/{k}"] = ref_grads[k].abs().max().numpy()
                out[f"train/grad_sum/{k}"] = ref_grads[k].double().sum().numpy()
                out[f"train/param1_sub/{k}"] = sub(ref_params[k])
            sqs = {k: L.optim.state[dict(L.net.named_parameters())[k]]["square_avg"] for k in NO.PARAM_KEYS}
            for k in NO.PARAM_KEYS:
                out[f"train/sqavg_sub/{k}"] = sub(sqs[k])
    for k in NO.PARAM_KEYS:
        out[f"train/param{n_steps}_sub/{k}"] = sub(L.net.state_dict()[k])
    out["train/priority_after3"] = L.replay_memory.memory["priority"].copy()
    out["train/seqprio_after3"] = L.replay_memory.memory["sequence_priority"].copy()
    out["train/config"] = np.array([M, B, BI, LN, NS, H])
    np.savez_compressed(os.path.join(GOLD, "network.npz"), **out)


# ------------------------------------------------------------------ hand-off dump golden
class _NoLock:
    """Stand-in for fasteners.InterProcessLock (not installed): single process, always free."""

    def __init__(self, path):
        self.path = path

    def acquire(self, blocking=True):
        return True

    def release(self):
        pass


def make_dump(ref_replay):
    """A dump file written by the REFERENCE's own ``ReplayMemory.save`` (fit + torch.save,
    replay_memory.py:121-152), rows appended through its ``add`` (float frames -> u8
    truncation): the wire format the learner's ``load`` must read."""
    sys.modules["fasteners"].InterProcessLock = _NoLock
    rows, H = 7, 256
    r = ref_replay.ReplayMemory(16, 4, 2)
    rng = np.random.RandomState(31)
    for i in range(rows):
        frame = rng.randint(0, 256, (4, 84, 84)).astype(np.float32) / np.float32(255)
        r.add(frame, (rng.randn(1, H) * 0.1).astype(np.float32), (rng.randn(1, H) * 0.1).astype(np.float32),
              (rng.randn(1, H) * 0.1).astype(np.float32), (rng.randn(1, H) * 0.1).astype(np.float32),
              int(rng.randint(6)), float(rng.randn()), bool(i == 4), 1, float(rng.rand() + 0.01))
    r.burn_in_length, r.learning_length, r.sequence_length = 2, 2, 4
    starts = np.array([0, 2, 3])
    r.update_sequence_priority(starts)
    r.memory["is_seq_start"][starts] = 1
    path = os.path.join(GOLD, "memory7.pt")
    if os.path.exists(path):
        os.remove(path)
    r.save(GOLD, 7)
    d = torch.load(path, weights_only=False)
    assert d["state"].shape == (rows, 4, 84, 84) and d["state"].dtype == np.uint8
    print("[dump] reference-written hand-off file:", {k: (v.dtype.name, v.shape) for k, v in d.items()})


# ------------------------------------------------------------------ actor goldens
def make_actor(ref_replay, ref_model):
    """The UNMODIFIED reference ``Actor`` (actor.py) stepped on the scripted environment of
    oracle/actor_oracle.py (``env`` is stubbed: gym/ALE are not installed).  Stores what the
    actor fed its replay (per step: action, reward, done, both nets' Q/h/c) and the replay it
    ended up with; asserts the oracle restatement reproduces that replay."""
    import contextlib
    import io
    from oracle import actor_oracle as AO
    env_mod = types.ModuleType("env")
    env_mod.PongEnv = None
    sys.modules["env"] = env_mod
    import actor as ref_actor  # noqa
    out = {}
    BI, LN, OV, NS = 3, 3, 3, 3
    cases = {"a": (1, 3, 11, (11, 9, 13)), "b": (2, 3, 12, (8, 2, 15, 7))}
    for name, (actor_id, n_actors, seed, lengths) in cases.items():
        ref_actor.PongEnv = lambda action_repeat, n_stacks, _s=seed, _l=lengths: AO.ScriptedEnv(_s, _l)
        torch.manual_seed(0)
        np.random.seed(100 + seed)
        A = ref_actor.Actor(actor_id, n_actors, {}, "cpu")
        A.burn_in_length, A.learning_length, A.overlap_length = BI, LN, OV
        A.sequence_length = BI + LN
        A.memory_save_interval = A.net_load_interval = 10 ** 9
        M = 64
        A.memory_size = M
        A.replay_memory = ref_replay.ReplayMemory(M, 4, NS)
        A.replay_memory.burn_in_length, A.replay_memory.learning_length = BI, LN
        A.replay_memory.sequence_length = BI + LN
        # make the two nets differ (the reference starts them equal): target = online * 1.01
        with torch.no_grad():
            for v in A.target_net.parameters():
                v.mul_(1.01)
        rec = []
        orig = A.select_action

        def spy(state, _o=orig, _rec=rec):
            r = _o(state)
            _rec.append(dict(state=np.array(state, copy=True), action=r[0], q=r[1].numpy().copy(), h=np.array(r[2]),
                             c=np.array(r[3]), tq=r[4].numpy().copy(), th=np.array(r[5]), tc=np.array(r[6])))
            return r
        A.select_action = spy
        orig_step = A.env.step

        def env_spy(action, _o=orig_step, _rec=rec):
            r = _o(action)
            _rec[-1]["reward"], _rec[-1]["done"] = r[1], r[2]
            return r
        A.env.step = env_spy
        n_steps = sum(lengths)
        with contextlib.redirect_stdout(io.StringIO()):
            for _ in range(n_steps):
                A.step()
        rm = A.replay_memory
        rows = rm.index
        # oracle restatement fed with the same per-step network outputs
        O = AO.ActorOracle(M, 256, NS, BI, LN, OV)
        for r in rec:
            O.observe(r["state"], r["action"], r["q"], r["h"], r["c"], r["tq"], r["th"], r["tc"], r["reward"], r["done"])
        assert O.replay.index == rows, (O.replay.index, rows)
        for k in RO.KEYS:
            a, b = O.replay.memory[k][:rows], rm.memory[k][:rows]
            if k in ("priority", "sequence_priority"):
                assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (name, k)
            else:
                assert np.array_equal(a, b), (name, k)
        print(f"[actor] case {name}: {n_steps} steps, {rows} rows stored; oracle restatement == reference replay (bit-exact)")
        # which step's frame each stored row holds (frames are regenerated from the env seed)
        src = []
        for i in range(rows):
            hit = [j for j, r in enumerate(rec) if np.array_equal((r["state"] * 255).astype(np.uint8), rm.memory["state"][i])]
            assert len(hit) == 1
            src.append(hit[0])
        out.update({f"{name}/config": np.array([actor_id, n_actors, seed, M, BI, LN, OV, NS]),
                    f"{name}/lengths": np.array(lengths), f"{name}/np_seed": 100 + seed,
                    f"{name}/action": np.array([r["action"] for r in rec]),
                    f"{name}/reward": np.array([r["reward"] for r in rec]),
                    f"{name}/done": np.array([r["done"] for r in rec]),
                    f"{name}/q": np.stack([r["q"].reshape(6) for r in rec]),
                    f"{name}/tq": np.stack([r["tq"].reshape(6) for r in rec]),
                    f"{name}/h_last": rec[-1]["h"].reshape(-1), f"{name}/c_last": rec[-1]["c"].reshape(-1),
                    f"{name}/rows": rows, f"{name}/state_src_step": np.array(src),
                    f"{name}/epsilon": A.epsilon})
        for k in RO.KEYS:
            if k != "state":
                out[f"{name}/mem/{k}"] = rm.memory[k][:rows].copy()
    np.savez_compressed(os.path.join(GOLD, "actor.npz"), **out)


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref_replay, ref_model, ref_learner = import_reference()
    torch.set_num_threads(8)
    only = set(sys.argv[1:])
    if not only or "priority" in only:
        make_priority(ref_replay)
    if not only or "sampling" in only:
        make_sampling(ref_replay)
    if not only or "network" in only:
        make_network(ref_replay, ref_model, ref_learner)
    if not only or "dump" in only:
        make_dump(ref_replay)
    if not only or "actor" in only:
        make_actor(ref_replay, ref_model)
    print("golden fixtures written to", GOLD)


if __name__ == "__main__":
    main()
