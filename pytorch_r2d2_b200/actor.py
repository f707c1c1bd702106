"""Actor-side inference of the hot path (/root/reference actor.py:96-106,144-157), batched.

The reference steps one environment per process at batch 1, evaluating the online AND the target
net every step (actor.py:99-100).  Here N actors share one forward per net through the same
CUDA kernels as the learner (torso -> input projection -> LSTM step -> dueling head, T=1),
each actor keeping its own (h, c) pair per net.  The environment loop, epsilon schedule and
n-step bookkeeping of actor.py stay outside the scope.
"""
from __future__ import annotations

import numpy as np
import torch

from .model import QNet


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

    def reset(self):
        self.net.reset()
        self.target_net.reset()

    @torch.no_grad()
    def forward_both(self, state):
        """state: [N,4,84,84] f32 in [0,1] on the device.  Returns (q, tq) [N,6]; hidden states
        advance inside the modules (model.py:48-57)."""
        q = self.net(state)
        tq = self.target_net(state)
        return q, tq

    @torch.no_grad()
    def select_action(self, state, rng=None):
        """Batched actor.py:96-106: epsilon-greedy over 6 actions."""
        q, tq = self.forward_both(state)
        greedy = q.argmax(1)
        rng = rng or np.random
        explore = torch.as_tensor(rng.random(self.n_actors) < self.epsilon, device=q.device)
        rand_a = torch.as_tensor(rng.randint(6, size=self.n_actors), device=q.device)
        action = torch.where(explore, rand_a, greedy)
        return action, q, tq

    def calc_priority(self, q_value, action, reward, next_q_value, target_next_q_value, done):
        """Vectorised actor.py:144-157 on host arrays ([N,6], [N], ...)."""
        q = np.asarray(q_value).reshape(-1, 6)[np.arange(len(action)), action]
        tn = np.asarray(target_next_q_value).reshape(-1, 6)
        na = np.asarray(next_q_value).reshape(-1, 6).argmax(-1)
        boot = reward + (self.gamma ** self.bootstrap_steps) * tn[np.arange(len(na)), na]
        target = np.where(np.asarray(done, bool), reward, boot)
        return (np.abs(q - target) + self.priority_epsilon) ** self.alpha
