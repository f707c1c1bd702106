"""Actor side of the hot path (/root/reference actor.py), batched on the device.

* ``BatchedActor``   -- N actors share one forward per net through the learner's kernels (torso ->
  input projection -> LSTM step -> dueling head, T=1, online AND target net every step like
  actor.py:99-100), each actor keeping its own (h, c) pair per net.
* ``ActorExperience`` -- what ``Actor.step`` does with the result (actor.py:74-94): n-step
  memory with the reference's quirks, ``calc_priority``, the append to the actor's local
  replay ring and, at the end of an episode, ``set_seq_start_index`` -- one CUDA kernel per
  environment step for all N actors (csrc/actor_kernels.cu), rows handed to the learner's replay
  device-to-device instead of through files.
* ``Actor``          -- the reference's class surface (``Actor(actor_id, n_actors, shared_dict,
  device)``, ``select_action`` returning the 7-tuple, ``step``, ``reset``, ``load_model``,
  ``calc_priority``, ``set_seq_start_index``) as a thin shim over the two with N = 1.

The environment itself (env.py: gym + ALE) is outside the scope; ``Actor`` takes any object
with ``reset()`` / ``step(action)`` as ``env``.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

from . import _lib
from .model import QNet
from .replay_memory import KEYS, ReplayMemory


def actor_process(actor_id, n_actors, shared_dict, device="cuda:0", env=None):
    Actor(actor_id, n_actors, shared_dict, device, env=env).run()


class BatchedActor:
    def __init__(self, n_actors, device="cuda:0", cell_size=256, net=None, target_net=None,
                 precision="fp32"):
        self.gamma = 0.99
        self.bootstrap_steps = 3
        self.alpha = 0.6
        self.priority_epsilon = 1e-6
        self.n_actors = n_actors
        self.device = torch.device(device)
        # epsilon_i = 0.4 ** (1 + 7 i / (N-1))  (actor.py:22; N=1 divides by zero there)
        i = np.arange(n_actors)
        self.epsilon = 0.4 ** (1 + i * 7 / max(n_actors - 1, 1))
        self.net = net if net is not None else QNet(device, cell_size, precision)
        self.target_net = target_net if target_net is not None else QNet(device, cell_size, precision)
        if target_net is None:
            self.target_net.load_state_dict(self.net.state_dict())
        # One step of both nets is ~50 small kernels: replayed from a CUDA graph over a fixed
        # workspace (nothing allocated per step, the weights' bf16 operand images refreshed inside
        # the graph, so weight updates need no invalidation).  False = enqueue kernel by kernel.
        self.use_cuda_graph = True
        self._g = None

    def _build_graph(self):
        from .model import alloc_forward_ws, net_forward
        N, dev = self.n_actors, self.device
        g = {"obs": torch.empty((N, 4, 84, 84), dtype=torch.float32, device=dev), "ws": [], "last": [None, None]}
        nets = (self.net, self.target_net)
        for net in nets:
            ws = alloc_forward_ws(net._weights(), 1, N, dev, want_gates=False, tc=net.precision == "bf16")
            ws["hs_ext"].zero_()
            ws["cs_ext"].zero_()
            g["ws"].append(ws)
        side = torch.cuda.Stream(device=dev)

        def both():
            main = torch.cuda.current_stream()
            side.wait_stream(main)
            for net, ws, stream in zip(nets, g["ws"], (main, side)):
                with torch.cuda.stream(stream):
                    net_forward(net._weights(), 1, N, ws["hs_ext"][0], ws["cs_ext"][0], frames=g["obs"],
                                want_gates=False, ws=ws, tc=net.precision == "bf16")
            main.wait_stream(side)

        g["obs"].zero_()
        saved = [(ws["hs_ext"].clone(), ws["cs_ext"].clone()) for ws in g["ws"]]
        s = torch.cuda.Stream(device=dev)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            both()                               # warm-up outside capture
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            both()
        for ws, (hs, cs) in zip(g["ws"], saved):
            ws["hs_ext"].copy_(hs)
            ws["cs_ext"].copy_(cs)
        g["graph"] = graph
        self._g = g
        return g

    def _graph_step(self, state):
        g = self._g or self._build_graph()
        for i, (net, ws) in enumerate(zip((self.net, self.target_net), g["ws"])):
            # module-held state (model.py:48-57): None -> zeros; anything the caller put there
            # (set_state, a partial reset) -> copied in; our own previous output -> carried over
            if net.hs is None:
                ws["hs_ext"][0].zero_()
                ws["cs_ext"][0].zero_()
            elif net.hs is g["last"][i]:
                ws["hs_ext"][0].copy_(ws["hs_ext"][1])
                ws["cs_ext"][0].copy_(ws["cs_ext"][1])
            else:
                ws["hs_ext"][0].copy_(torch.as_tensor(net.hs).to(self.device))
                ws["cs_ext"][0].copy_(torch.as_tensor(net.cs).to(self.device))
        g["obs"].copy_(state.reshape(g["obs"].shape), non_blocking=True)
        g["graph"].replay()
        for i, (net, ws) in enumerate(zip((self.net, self.target_net), g["ws"])):
            net.hs, net.cs = ws["hs_ext"][1], ws["cs_ext"][1]
            g["last"][i] = net.hs
        return g["ws"][0]["q"], g["ws"][1]["q"]

    def reset(self, actors=None):
        """Zero recurrent state (actor.py:112-113) for all actors, or for the given ones."""
        for net in (self.net, self.target_net):
            if actors is None or net.hs is None:
                net.reset()
            else:
                idx = torch.as_tensor(actors, device=self.device, dtype=torch.int64)
                net.hs = net.hs.clone()
                net.cs = net.cs.clone()
                net.hs[idx] = 0
                net.cs[idx] = 0

    @torch.no_grad()
    def forward_both(self, state):
        """state: [N,4,84,84] f32 in [0,1] on the device.  Returns (q, tq) [N,6]; hidden states
        advance inside the modules (model.py:48-57).  With ``use_cuda_graph`` the returned tensors
        (and ``net.hs`` / ``net.cs``) live in the fixed workspace: valid until the next step."""
        if self.use_cuda_graph and state.shape[0] == self.n_actors:
            return self._graph_step(state)
        q = self.net(state)
        tq = self.target_net(state)
        return q, tq

    @torch.no_grad()
    def select_action(self, state, rng=None):
        """Batched actor.py:96-106.  Returns ``(action, q, h, c, target_q, target_h, target_c)``:
        device tensors [N], [N,6], [N,H], ... (the reference's 7-tuple with a leading actor axis).
        Epsilon-greedy draws follow the reference's order per actor -- ``random()``, then
        ``randint(6)`` only when exploring -- from `rng`: one numpy RandomState-like object for
        all actors (default: numpy's global one) or a list with one per actor."""
        q, tq = self.forward_both(state)
        h, c, th, tc = self.net.hs, self.net.cs, self.target_net.hs, self.target_net.cs
        greedy = q.argmax(1).cpu().numpy()
        action = np.empty(self.n_actors, np.int64)
        for a in range(self.n_actors):
            r = rng[a] if isinstance(rng, (list, tuple)) else (rng or np.random)
            action[a] = r.randint(6) if r.random() < self.epsilon[a] else greedy[a]
        return torch.from_numpy(action).to(self.device), q, h, c, tq, th, tc

    def calc_priority(self, q_value, action, reward, next_q_value, target_next_q_value, done):
        """Vectorised actor.py:144-157 on host arrays ([N,6], [N], ...), float32 arithmetic like
        numpy 2 gives the reference."""
        f = np.float32
        a = np.asarray(action).reshape(-1)
        q = np.asarray(q_value, f).reshape(-1, 6)[np.arange(len(a)), a]
        tn = np.asarray(target_next_q_value, f).reshape(-1, 6)
        na = np.asarray(next_q_value, f).reshape(-1, 6).argmax(-1)
        r = np.asarray(reward, np.float64).astype(f)
        boot = r + f(self.gamma ** self.bootstrap_steps) * tn[np.arange(len(na)), na]
        target = np.where(np.asarray(done, bool), r, boot).astype(f)
        return (np.abs(q - target) + f(self.priority_epsilon)) ** f(self.alpha)


class _ActorBuffers(ctypes.Structure):
    """Mirror of r2d2_actor_buffers (include/r2d2_b200.h)."""
    _fields_ = [(k, ctypes.c_void_p) for k in (
        "ns_state", "ns_hc", "ns_thc", "ns_q", "ns_action", "ns_reward", "counters",
        "state", "hs_cs", "target_hs_cs", "action", "reward", "done", "stack_count", "priority",
        "sequence_priority", "is_seq_start", "index", "ep_start")]


class ActorExperience:
    """Device-side experience construction for N actors (csrc/actor_kernels.cu).

    Every actor owns a local ring of `capacity` rows with the ten replay keys (tensors shaped
    [N, capacity, ...]); ``observe`` is one kernel launch per environment step; ``flush_to``
    appends the rows of finished episodes to a learner's ``ReplayMemory`` device-to-device (the
    reference's actors write their local replay to a file every five episodes instead,
    actor.py:125-129).  ``rings`` lets a caller supply the ring tensors (the single-actor
    ``Actor`` shares them with its ``ReplayMemory``)."""

    _DT = ReplayMemory._DT

    def __init__(self, n_actors, device="cuda:0", cell_size=256, n_step=3, burn_in_length=10,
                 learning_length=10, overlap_length=10, capacity=2048, gamma=0.99, alpha=0.6,
                 priority_epsilon=1e-6, eta=0.9, rings=None):
        if not torch.cuda.is_available():
            raise RuntimeError("pytorch_r2d2_b200.ActorExperience needs a CUDA device (no CPU fallback)")
        _lib.load()
        self.N, self.n, self.H, self.C = n_actors, n_step, cell_size, capacity
        self.burn_in_length, self.learning_length = burn_in_length, learning_length
        self.overlap_length = overlap_length
        self.gamma, self.alpha, self.priority_epsilon, self.eta = gamma, alpha, priority_epsilon, eta
        self.device = dev = torch.device(device)
        N, n, H, C = self.N, self.n, self.H, self.C
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)  # noqa: E731
        self.ns = dict(ns_state=z((N, n, 28224), torch.uint8), ns_hc=z((N, n, 2 * H), torch.float32),
                       ns_thc=z((N, n, 2 * H), torch.float32), ns_q=z((N, n, 6), torch.float32),
                       ns_action=z((N, n), torch.int32), ns_reward=z((N, n), torch.float64),
                       counters=z((N, 2), torch.int32))
        if rings is None:
            rings = dict(state=z((N, C, 4, 84, 84), torch.uint8), hs_cs=z((N, C, 2 * H), torch.float32),
                         target_hs_cs=z((N, C, 2 * H), torch.float32), action=z((N, C, 1), torch.int8),
                         reward=z((N, C, 1), torch.float32), done=z((N, C, 1), torch.float32),
                         stack_count=z((N, C), torch.int8), priority=z((N, C), torch.float32),
                         sequence_priority=z((N, C), torch.float32), is_seq_start=z((N, C), torch.uint8))
        self.rings = rings
        self.index = z((N,), torch.int64)
        self.ep_start = z((N,), torch.int64)
        self._buf = _ActorBuffers(**{k: v.data_ptr() for k, v in {**self.ns, **self.rings}.items()},
                                  index=self.index.data_ptr(), ep_start=self.ep_start.data_ptr())
        # gamma**i as Python floats, exactly what NStepMemory.get multiplies with (replay_memory.py:51)
        self._gpow = torch.tensor([gamma ** i for i in range(n)], dtype=torch.float64, device=dev)
        self._steps_since_flush = 0

    @property
    def sequence_length(self):
        return self.burn_in_length + self.learning_length

    def observe(self, obs, action, q, h, c, tq, th, tc, reward, done):
        """One environment step of every actor (actor.py:74-94 after ``env.step``).
        obs [N,4,84,84] f32 in [0,1] (the state the action was chosen in); action [N]; q / tq
        [N,6], h c th tc [N,H] as returned by ``BatchedActor.select_action``; reward [N] floats,
        done [N] bools from the environments."""
        if self._steps_since_flush + 1 > self.C:
            raise RuntimeError(f"actor-local ring would wrap ({self.C} rows): flush_to() the learner's replay first")
        dev, N = self.device, self.N
        t = lambda x, dt: torch.as_tensor(np.asarray(x) if not torch.is_tensor(x) else x).to(  # noqa: E731
            device=dev, dtype=dt, non_blocking=True).contiguous()
        obs = t(obs, torch.float32).reshape(N, 28224)
        args = [t(action, torch.int64).reshape(N), t(q, torch.float32).reshape(N, 6),
                t(h, torch.float32).reshape(N, self.H), t(c, torch.float32).reshape(N, self.H),
                t(tq, torch.float32).reshape(N, 6), t(th, torch.float32).reshape(N, self.H),
                t(tc, torch.float32).reshape(N, self.H), t(reward, torch.float64).reshape(N),
                t(done, torch.uint8).reshape(N)]
        self._keep = (obs, args)       # the kernel reads them asynchronously
        f32 = lambda v: float(np.float32(v))  # noqa: E731
        _lib.call("r2d2_actor_observe", ctypes.byref(self._buf), N, self.n, self.H, self.C,
                  self.sequence_length, self.overlap_length, _lib.ptr(obs, "f32"),
                  *[ctypes.c_void_p(x.data_ptr()) for x in args], _lib.ptr(self._gpow, "f64"),
                  f32(self.gamma ** self.n), f32(self.priority_epsilon), f32(self.alpha),
                  f32(self.eta), f32(1 - self.eta), _lib.stream_ptr())
        self._steps_since_flush += 1

    def rows(self):
        """Rows written per actor since the last flush (host ints; synchronises)."""
        return self.index.cpu().tolist()

    def dump(self, actor):
        """``{key: ndarray[:rows]}`` of one actor's local ring (the reference's dump payload)."""
        n = int(self.index[actor].item())
        return {k: self.rings[k][actor, :n].cpu().numpy() for k in KEYS}

    def flush_to(self, replay: ReplayMemory, actors=None, finished_only=True):
        """Append the local rows of `actors` (default all) to the learner's replay, device to
        device, and restart their rings.  With `finished_only` the rows of an episode still in
        progress stay behind (they have no sequence starts yet) and move to the ring's front."""
        idx, eps = self.index.cpu().tolist(), self.ep_start.cpu().tolist()
        for a in (range(self.N) if actors is None else actors):
            n = eps[a] if finished_only else idx[a]
            if n > 0:
                replay.extend({k: self.rings[k][a, :n] for k in KEYS})
            rest = idx[a] - n
            if rest > 0:
                for k in KEYS:
                    self.rings[k][a, :rest] = self.rings[k][a, n:idx[a]].clone()
            for k in ("sequence_priority", "is_seq_start"):
                self.rings[k][a, rest:idx[a]] = 0
            self.index[a] = rest
            self.ep_start[a] = 0
        self._steps_since_flush = max(self.index.cpu().tolist())


class Actor:
    """The reference's ``Actor`` surface (actor.py:18-167) on the device path, N = 1.

    Same constructor arguments and attributes; ``env`` (any object with ``reset()`` ->
    state[4,84,84] floats in [0,1] and ``step(action)`` -> (next_state, reward, done, info))
    replaces the hard-wired ``PongEnv`` (gym + ALE are not part of this library).
    ``replay_memory`` is a device ``ReplayMemory`` whose ring the experience kernel writes;
    its ``index`` is refreshed from the device at the end of every episode."""

    def __init__(self, actor_id, n_actors, shared_dict, device="cuda:0", *, env=None, cell_size=256,
                 memory_size=50000, precision="fp32"):
        # params (actor.py:20-27)
        self.gamma = 0.99
        self.epsilon = 0.4 ** (1 + actor_id * 7 / (n_actors - 1))     # n_actors = 1 raises, as in the reference
        self.bootstrap_steps = 3
        self.alpha = 0.6
        self.priority_epsilon = 1e-6
        self.device = device
        self.actor_id = actor_id
        self.memory_path = os.path.join("./", "logs", "memory")
        # memory (actor.py:33-47)
        self.memory_size = memory_size
        self.batch_size = 32
        self.action_repeat = 4
        self.n_stacks = 4
        self.burn_in_length = 10
        self.learning_length = 10
        self.overlap_length = 10
        self.eta = 0.9
        self.sequence_length = self.burn_in_length + self.learning_length
        self.stack_count = self.n_stacks // self.action_repeat
        self.memory_save_interval = 5
        self.episode_start_index = 0
        self.cell_size = cell_size
        # net (actor.py:50-54): online net first, then the target net, like the reference's RNG order
        self.shared_dict = shared_dict
        self.net_load_interval = 5
        self._batched = BatchedActor(1, device, cell_size, precision=precision)
        self.net, self.target_net = self._batched.net, self._batched.target_net
        self._experience = None
        self.replay_memory = None
        self._new_local_replay()
        # env (actor.py:57-62)
        self.env = env
        self.episode_reward = 0
        self.n_episodes = 0
        self.n_steps = 0
        self.memory_count = 0
        self.state = self.env.reset() if env is not None else None

    def _new_local_replay(self):
        rm = ReplayMemory(self.memory_size, self.batch_size, self.bootstrap_steps, cell_size=self.cell_size,
                          device=self.device)
        self.replay_memory = rm
        self._experience = None          # built on first use, so that lengths set after construction count

    def _exp(self):
        rm = self.replay_memory
        e = self._experience
        shape = (self.bootstrap_steps, self.burn_in_length, self.learning_length, self.overlap_length)
        if e is None or e._shape != shape:
            rm.burn_in_length, rm.learning_length = self.burn_in_length, self.learning_length
            rm.sequence_length = self.sequence_length = self.burn_in_length + self.learning_length
            rings = {k: rm._raw(k).unsqueeze(0) for k in KEYS}
            old = e
            e = ActorExperience(1, self.device, self.cell_size, self.bootstrap_steps, self.burn_in_length,
                                self.learning_length, self.overlap_length, capacity=rm.memory_size,
                                gamma=self.gamma, alpha=self.alpha, priority_epsilon=self.priority_epsilon,
                                eta=self.eta, rings=rings)
            e._shape = shape
            if old is not None:
                e.index.copy_(old.index)
                e.ep_start.copy_(old.ep_start)
            self._experience = e
        return e

    def run(self):
        while True:
            self.step()

    def step(self):
        """actor.py:68-94."""
        if self.env is None:
            raise RuntimeError("Actor.step needs an environment: pass env=... (env.py's PongEnv needs gym + ALE)")
        state = self.state
        action, q_value, h, c, target_q_value, target_h, target_c = self.select_action(state)
        next_state, reward, done, _ = self.env.step(action)
        self.episode_reward += reward
        self.n_steps += 1
        self._exp().observe(np.asarray(state, np.float32)[-self.action_repeat:], [action], q_value, h, c,
                            target_q_value, target_h, target_c, [float(reward)], [bool(done)])
        self.state = np.array(next_state, copy=True)
        if done:
            self.reset()

    def select_action(self, state):
        """actor.py:96-106: ``(action, q_value, h, c, target_q_value, target_h, target_c)`` --
        action a Python int, the Q-values device tensors [1,6], the states numpy [1,H]."""
        state = torch.as_tensor(np.asarray(state, np.float32)).unsqueeze(0).to(self.device)
        with torch.no_grad():
            q_value, h, c = self.net(state, True)
            target_q_value, target_h, target_c = self.target_net(state, True)
        if np.random.random() < self.epsilon:
            action = np.random.randint(6)
        else:
            action = q_value.argmax().item()
        return action, q_value, h, c, target_q_value, target_h, target_c

    def reset(self):
        """actor.py:108-135."""
        print("episodes:", self.n_episodes, "actor_id:", self.actor_id, "return:", self.episode_reward)
        self.net.reset()
        self.target_net.reset()
        # set_seq_start_index already ran on the device when the episode's last step was observed
        self.replay_memory.index = int(self._exp().index[0].item())
        self.replay_memory._tree_dirty = True
        self.state = self.env.reset()
        self.episode_start_index = self.replay_memory.index
        self.episode_reward = 0
        self.n_episodes += 1
        self.n_steps = 0
        self.memory_count = 0
        self.stack_count = self.n_stacks // self.action_repeat
        if self.n_episodes % self.memory_save_interval == 0:
            self.replay_memory.save(self.memory_path, self.actor_id)
            self._new_local_replay()
            self.episode_start_index = 0
        if self.n_episodes % self.net_load_interval == 0:
            self.load_model()

    def load_model(self):
        """actor.py:137-142."""
        try:
            self.net.load_state_dict(self.shared_dict["net_state"])
            self.target_net.load_state_dict(self.shared_dict["target_net_state"])
        except Exception:       # noqa: BLE001 -- as the reference: report and keep the old weights
            print("load error")

    def calc_priority(self, q_value, action, reward, next_q_value, target_next_q_value, done):
        """actor.py:144-157 for one transition (host numpy; the device path computes the same
        inside ``ActorExperience.observe``)."""
        return self._batched.calc_priority(np.asarray(q_value).reshape(1, 6), [action], [reward],
                                           np.asarray(next_q_value).reshape(1, 6),
                                           np.asarray(target_next_q_value).reshape(1, 6), [done])[0]

    def set_seq_start_index(self):
        """actor.py:159-167 through the replay's own entry points (the per-step kernel does this at
        the end of every episode; this is the explicit form for callers that drive it by hand)."""
        last_index = self.replay_memory.index
        seq_start_index = [i for i in range(self.episode_start_index, last_index - self.sequence_length,
                                            self.overlap_length)]
        seq_start_index.append(last_index - self.sequence_length)
        seq_start_index = np.array([i for i in seq_start_index if i >= 0])
        self.replay_memory.update_sequence_priority(seq_start_index)
        self.replay_memory.memory["is_seq_start"][torch.from_numpy(seq_start_index).to(self.device)] = 1
