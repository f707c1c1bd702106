"""GPU: the actor side -- the reference-compatible ``Actor``
against the replay the UNMODIFIED reference Actor built on the same scripted environment
(tests/golden/actor.npz, oracle/make_golden.py actor), the batched experience kernel against
the oracle restatement, and batched inference at BASELINE config 5's shape."""
import contextlib
import io
import os

import numpy as np
import pytest
import torch

from oracle import actor_oracle as AO
from oracle import net_oracle as NO
from oracle import replay_oracle as RO
from util import gpu_mem_np, rel_err

pytestmark = pytest.mark.gpu
KEYS = RO.KEYS


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "actor.npz"))


@pytest.mark.parametrize("name", ["a", "b"])
def test_actor_step_vs_reference_golden(cuda_device, gold, name):
    """``Actor.step`` driven exactly like the reference's (same seeds, same scripted env): the
    local replay it builds == the reference actor's -- frames, actions, n-step rewards, done
    flags, sequence starts bit-exact; hidden states / priorities in the fp32 class."""
    from pytorch_r2d2_b200 import Actor
    g = gold
    actor_id, n_actors, seed, M, BI, LN, OV, NS = [int(v) for v in g[f"{name}/config"]]
    lengths = tuple(int(x) for x in g[f"{name}/lengths"])
    torch.manual_seed(0)
    np.random.seed(int(g[f"{name}/np_seed"]))
    A = Actor(actor_id, n_actors, {}, "cuda:0", env=AO.ScriptedEnv(seed, lengths), memory_size=M)
    assert A.epsilon == float(g[f"{name}/epsilon"])
    A.burn_in_length, A.learning_length, A.overlap_length = BI, LN, OV
    A.sequence_length = BI + LN
    A.memory_save_interval = A.net_load_interval = 10 ** 9
    with torch.no_grad():
        A.target_net.flat_parameters().mul_(1.01)           # as make_golden: nets differ
    n_steps = len(g[f"{name}/action"])
    actions, qs = [], []
    orig = A.select_action

    def spy(state):
        r = orig(state)
        actions.append(r[0])
        qs.append(r[1].cpu().numpy().reshape(6))
        assert isinstance(r[0], int) and r[1].shape == (1, 6) and isinstance(r[2], np.ndarray) and r[2].shape == (1, 256)
        assert len(r) == 7 and r[4].shape == (1, 6) and r[6].shape == (1, 256)
        return r
    A.select_action = spy
    with contextlib.redirect_stdout(io.StringIO()):
        for _ in range(n_steps):
            A.step()
    assert actions == g[f"{name}/action"].tolist()
    assert rel_err(np.stack(qs), g[f"{name}/q"]) < 1e-5
    rows = int(g[f"{name}/rows"])
    assert A.replay_memory.index == rows
    d = A.replay_memory.dump()
    for k in ("action", "reward", "done", "stack_count", "is_seq_start"):
        assert np.array_equal(d[k], g[f"{name}/mem/{k}"]), k
    for k in ("hs_cs", "target_hs_cs"):
        assert rel_err(d[k], g[f"{name}/mem/{k}"]) < 1e-5, k
    for k in ("priority", "sequence_priority"):
        assert np.allclose(d[k], g[f"{name}/mem/{k}"], rtol=1e-4, atol=1e-7), k
    # frames: regenerate the scripted episodes, row i holds the frame of step state_src_step[i]
    env = AO.ScriptedEnv(seed, lengths)
    state, frames = env.reset(), []
    for t in range(n_steps):
        frames.append(state)
        nxt, _, done, _ = env.step(0)
        state = env.reset() if done else nxt
    for i, s in enumerate(g[f"{name}/state_src_step"]):
        assert np.array_equal(d["state"][i], (frames[s] * 255).astype(np.uint8)), i
    # sequence priorities are bit-exact given the per-step priorities the device wrote
    chk = d["sequence_priority"].copy()
    starts = np.nonzero(d["is_seq_start"])[0]
    sp = np.zeros(M, np.float32)
    pr = np.zeros(M, np.float32)
    pr[:rows] = d["priority"]
    RO.update_sequence_priority(pr, sp, np.zeros(M, np.uint8), starts, BI + LN, 0.9, False)
    assert np.array_equal(sp[:rows].view(np.uint32), chk.view(np.uint32))


@pytest.mark.parametrize("n,T,OV", [(3, 6, 3), (5, 20, 10), (1, 4, 2)])
def test_batched_experience_kernel_vs_oracle(cuda_device, n, T, OV):
    """N actors, one kernel launch per environment step, network outputs random (the kernel is
    independent of the network): every actor's local ring == its oracle twin, then the rows
    reach a learner's replay device-to-device exactly as a reference ``extend`` of the dumps."""
    from pytorch_r2d2_b200 import ActorExperience, ReplayMemory
    N, H, C = 5, 256, 160
    BI = T // 2
    rng = np.random.RandomState(n)
    E = ActorExperience(N, "cuda:0", H, n, BI, T - BI, OV, capacity=C)
    twins = [AO.ActorOracle(C, H, n, BI, T - BI, OV) for _ in range(N)]
    lengths = [[T + 9, 2, 1, T + 3, 7, 30], [T + 20, 3, 40], [T + 2, T + 2, 11, 1, 1, 25],
               [60, 4, 9, 9, 9], [T + 1, 35, 2, 8]]
    t_in_ep, ep = [0] * N, [0] * N
    for step in range(70):
        obs = rng.randint(0, 256, (N, 4, 84, 84)).astype(np.float32) / np.float32(255)
        q, tq = rng.randn(N, 6).astype(np.float32), rng.randn(N, 6).astype(np.float32)
        h, c, th, tc = (rng.randn(N, H).astype(np.float32) for _ in range(4))
        action = rng.randint(6, size=N)
        reward = rng.choice([-1.0, 0.0, 0.5, 1.0], N)
        done = np.zeros(N, bool)
        for a in range(N):
            t_in_ep[a] += 1
            if t_in_ep[a] >= lengths[a][ep[a] % len(lengths[a])]:
                done[a], t_in_ep[a], ep[a] = True, 0, ep[a] + 1
        E.observe(torch.from_numpy(obs).cuda(), action, q, h, c, tq, th, tc, reward, done)
        for a in range(N):
            twins[a].observe(obs[a], int(action[a]), q[a][None], h[a][None], c[a][None], tq[a][None],
                             th[a][None], tc[a][None], float(reward[a]), bool(done[a]))
    torch.cuda.synchronize()
    assert E.rows() == [tw.replay.index for tw in twins]
    assert E.ep_start.cpu().tolist() == [tw.episode_start_index for tw in twins]
    for a in range(N):
        d, m = E.dump(a), twins[a].replay.memory
        rows = twins[a].replay.index
        for k in ("state", "hs_cs", "target_hs_cs", "action", "reward", "done", "stack_count", "is_seq_start"):
            assert np.array_equal(d[k], m[k][:rows]), (a, k)
        assert np.allclose(d["priority"], m["priority"][:rows], rtol=2e-6, atol=0), a
        assert np.allclose(d["sequence_priority"], m["sequence_priority"][:rows], rtol=2e-6, atol=0), a
        sp, pr = np.zeros(C, np.float32), np.zeros(C, np.float32)
        pr[:rows] = d["priority"]
        RO.update_sequence_priority(pr, sp, np.zeros(C, np.uint8), np.nonzero(d["is_seq_start"])[0], T, 0.9, False)
        assert np.array_equal(sp[:rows].view(np.uint32), d["sequence_priority"].view(np.uint32)), a
    # hand-off: finished episodes go to the learner's ring, the episode in progress stays local
    learner_rm = ReplayMemory(1024, 4, n, cell_size=H, device="cuda:0")
    twin_rm = RO.ReplayOracle(1024, H, n, BI, T - BI)
    dumps = [E.dump(a) for a in range(N)]
    eps = E.ep_start.cpu().tolist()
    E.flush_to(learner_rm)
    for a in range(N):
        if eps[a]:
            twin_rm.extend({k: v[:eps[a]] for k, v in dumps[a].items()})
    assert learner_rm.index == twin_rm.index
    for k in KEYS:
        assert np.array_equal(gpu_mem_np(learner_rm, k), twin_rm.memory[k]), k
    assert E.rows() == [tw.replay.index - e for tw, e in zip(twins, eps)]
    for a in range(N):
        rest = E.rows()[a]
        assert np.array_equal(E.dump(a)["state"], dumps[a]["state"][eps[a]:eps[a] + rest])


def test_calc_priority_host_vs_oracle(cuda_device):
    """actor.py:144-157, both branches, against the oracle restatement (itself bit-exact with
    the reference actor, oracle/make_golden.py)."""
    from pytorch_r2d2_b200 import BatchedActor
    act = BatchedActor(7, "cuda:0", 256)
    rng = np.random.RandomState(3)
    q, nq, tq = (rng.randn(7, 6).astype(np.float32) for _ in range(3))
    action, reward = rng.randint(6, size=7), rng.randn(7)
    done = np.array([0, 1, 0, 0, 1, 0, 1], bool)
    got = act.calc_priority(q, action, reward, nq, tq, done)
    want = [AO.calc_priority(q[i][None], int(action[i]), float(reward[i]), nq[i][None], tq[i][None], bool(done[i]))
            for i in range(7)]
    assert got.dtype == np.float32 and np.allclose(got, want, rtol=2e-7, atol=0)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_batched_inference_config5_shape(cuda_device, precision):
    """BASELINE config 5: 256 actors, LSTM 512, both nets, three consecutive steps with the
    recurrent state carried per actor == the oracle's QNet.forward on the same batch."""
    from pytorch_r2d2_b200 import BatchedActor
    N, H = 256, 512
    torch.manual_seed(11)
    act = BatchedActor(N, "cuda:0", H, precision=precision)
    with torch.no_grad():
        act.target_net.flat_parameters().mul_(0.97)
    p = {k: v.cpu().clone() for k, v in act.net.state_dict().items()}
    pt = {k: v.cpu().clone() for k, v in act.target_net.state_dict().items()}
    rng = np.random.RandomState(4)
    h = c = th = tc = None
    for step in range(3):
        x = torch.from_numpy(rng.randint(0, 256, (N, 4, 84, 84)).astype(np.float32) / np.float32(255))
        with torch.no_grad():
            q_ref, _, _, (h, c) = NO.qnet_forward(p, x, h, c)
            tq_ref, _, _, (th, tc) = NO.qnet_forward(pt, x, th, tc)
        action, q, hh, cc, tq, thh, tcc = act.select_action(x.cuda(), rng=np.random.RandomState(step))
        errs = [rel_err(q, q_ref), rel_err(tq, tq_ref), rel_err(hh, h), rel_err(cc, c), rel_err(thh, th), rel_err(tcc, tc)]
        assert max(errs) < 1e-5, (step, errs)
        assert action.shape == (N,) and hh.shape == (N, H)
