"""Oracle (CPU restatement, numpy) of the reference's actor-side experience construction:
``NStepMemory`` (replay_memory.py:12-55), ``Actor.step`` / ``calc_priority`` /
``set_seq_start_index`` (actor.py:68-94,144-167), and a scripted environment that stands in for
``PongEnv`` (env.py needs gym + ALE, which the image does not have).

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.  Pinned by ``oracle/make_golden.py actor``:
the UNMODIFIED reference ``Actor`` is run on the scripted environment (only ``env`` and
``fasteners`` are stubbed) and this restatement must reproduce its replay contents.

Reference behaviour worth knowing (all reproduced):
  * the n-step deques have ``maxlen=n`` and the first ``get`` happens at step n+1, after the
    append has already pushed the first transition out: the first transition of every episode
    is never stored (actor.py:77-81, replay_memory.py:17-25);
  * ``get`` pops every deque except ``reward`` (replay_memory.py:41-52), so the reward deque
    always holds the last <= n rewards and every transition flushed at the end of an episode
    gets the SAME discounted sum over them (actor.py:88-93);
  * the priority of a stored transition uses the Q-values of the CURRENT step as "next" values
    (actor.py:82), and ``done`` of the current step;
  * sequence starts: every ``overlap_length`` rows from the episode's first row while
    ``< last - T``, plus a forced start at ``last - T`` (actor.py:159-167).  If the local ring
    holds fewer than T rows at the end of an episode the reference indexes with a negative
    slot (and raises on an empty slice); implementations may skip that forced start -- the
    fixtures never get there.
"""
from __future__ import annotations

from collections import deque

import numpy as np

from . import replay_oracle as RO


class ScriptedEnv:
    """Deterministic stand-in for ``PongEnv`` (env.py:8-37): 4x84x84 float frames that are exact
    u8/255 values, rewards in {-1, 0, 1}, episode ends after scripted lengths; independent of
    the actions taken."""

    def __init__(self, seed=0, episode_lengths=(11, 4, 9)):
        self.rng = np.random.RandomState(seed)
        self.lengths = list(episode_lengths)
        self.episode = -1
        self.t = 0

    def _frame(self):
        return self.rng.randint(0, 256, (4, 84, 84)).astype(np.float32) / np.float32(255)

    def reset(self):
        self.episode += 1
        self.t = 0
        return self._frame()

    def step(self, action):
        self.t += 1
        reward = float(self.rng.choice([-1.0, 0.0, 0.0, 1.0]))
        done = self.t >= self.lengths[self.episode % len(self.lengths)]
        return self._frame(), reward, bool(done), {}


class NStepOracle:
    """replay_memory.py:12-55: nine deques of maxlen n; ``get`` pops all but ``reward``."""

    def __init__(self, n=3, gamma=0.99):
        self.n, self.gamma = n, gamma
        self.items = deque(maxlen=n)     # (q, state, h, c, th, tc, action, stack_count)
        self.reward = deque(maxlen=n)

    @property
    def size(self):
        return len(self.items)

    def add(self, q, state, h, c, th, tc, action, reward, stack_count):
        self.items.append((q, state, h, c, th, tc, action, stack_count))
        self.reward.append(reward)

    def get(self):
        q, state, h, c, th, tc, action, stack_count = self.items.popleft()
        reward = sum([self.gamma ** i * r for i, r in enumerate(self.reward)])   # python floats (f64)
        return q, state, h, c, th, tc, action, reward, stack_count


def calc_priority(q_value, action, reward, next_q_value, target_next_q_value, done,
                  gamma=0.99, n=3, eps=1e-6, alpha=0.6):
    """actor.py:144-157.  q arrays are float32 ([1,6]); under numpy 2 the Python-float constants
    are weak scalars, so every operation rounds to float32.  Shapes are kept as the reference
    has them (a (1,)-array when bootstrapping, a scalar when done): numpy's array and scalar
    ``power`` paths need not agree in the last bit."""
    q = np.asarray(q_value, np.float32).reshape(-1)[action]
    tn = np.asarray(target_next_q_value, np.float32).reshape(-1)
    if done:
        target = reward
    else:
        na = np.asarray(next_q_value, np.float32).reshape(1, -1).argmax(-1)     # array([k])
        target = reward + (gamma ** n) * tn[na]
    p = np.abs(q - target) + eps
    p = p ** alpha
    return np.float32(np.asarray(p).reshape(-1)[0])


class ActorOracle:
    """One actor's experience construction (actor.py:68-94,108-135,159-167) into a local
    ``ReplayOracle``; the network outputs are inputs (q, h, c of both nets for the current
    state), so this is independent of the network implementation."""

    def __init__(self, memory_size=50000, cell=256, n=3, burn_in=10, learn=10, overlap=10,
                 gamma=0.99, alpha=0.6, eps=1e-6):
        self.n, self.gamma, self.alpha, self.eps = n, gamma, alpha, eps
        self.overlap = overlap
        self.replay = RO.ReplayOracle(memory_size, cell, n, burn_in, learn)
        self.nstep = NStepOracle(n, gamma)
        self.n_steps = 0
        self.episode_start_index = 0
        self.stack_count = 1

    def observe(self, state, action, q, h, c, tq, th, tc, reward, done):
        """What ``Actor.step`` does after ``env.step`` returned (actor.py:74-94)."""
        self.n_steps += 1
        self.nstep.add(q, state, h, c, th, tc, action, reward, self.stack_count)
        if self.n_steps > self.n:
            self._emit(q, tq, done)
        if done:
            while self.nstep.size > 0:
                self._emit(q, tq, done)
            self.end_episode()

    def _emit(self, q_now, tq_now, done):
        pq, state, h, c, th, tc, action, reward, sc = self.nstep.get()
        prio = calc_priority(pq, action, reward, q_now, tq_now, done, self.gamma, self.n, self.eps, self.alpha)
        self.replay.add(state, h, c, th, tc, action, reward, done, sc, prio)

    def end_episode(self):
        """``reset`` (actor.py:108-122): mark the episode's sequence starts, start a new episode."""
        last, T = self.replay.index, self.replay.T
        starts = [i for i in range(self.episode_start_index, last - T, self.overlap)]
        starts.append(last - T)
        starts = np.array(starts)
        self.replay.update_sequence_priority(starts)
        self.replay.memory["is_seq_start"][starts] = 1
        self.episode_start_index = self.replay.index
        self.n_steps = 0
        self.nstep = NStepOracle(self.n, self.gamma)
