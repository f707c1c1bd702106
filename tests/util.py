"""Shared helpers for the tests: synthetic replay contents mirrored into the oracle and the
GPU replay, error metrics."""
import numpy as np
import torch

from oracle import replay_oracle as RO


def rel_err(a, b):
    a = torch.as_tensor(a).detach().to(torch.float64).cpu().reshape(-1)
    b = torch.as_tensor(b).detach().to(torch.float64).cpu().reshape(-1)
    den = b.abs().max().clamp_min(1e-30)
    return float(((a - b).abs().max() / den).detach())


def fill_small(mem, M, H, seed):
    """Same generator as oracle/make_golden.py:fill_small."""
    rng = np.random.RandomState(seed)
    N = M
    mem["state"][:N] = rng.randint(0, 256, (N, 4, 84, 84)).astype(np.uint8)
    mem["hs_cs"][:N] = (rng.randn(N, 2 * H) * 0.1).astype(np.float32)
    mem["target_hs_cs"][:N] = (rng.randn(N, 2 * H) * 0.1).astype(np.float32)
    mem["action"][:N, 0] = rng.randint(6, size=N).astype(np.int8)
    mem["reward"][:N, 0] = rng.randn(N).astype(np.float32)
    mem["done"][:N, 0] = (np.arange(N) % 13 == 12).astype(np.float32)
    mem["stack_count"][:N] = 1
    mem["priority"][:N] = (rng.rand(N) + 0.01).astype(np.float32)


def make_oracle_replay(M, H, n_step, bi, ln, seed, start_stride=3):
    rep = RO.ReplayOracle(M, H, n_step, bi, ln)
    fill_small(rep.memory, M, H, seed)
    rep.index = M
    starts = np.arange(0, M - (bi + ln) - n_step, start_stride)
    rep.update_sequence_priority(starts)
    rep.memory["is_seq_start"][starts] = 1
    return rep, starts


def mirror_to_gpu(rep, batch_size, device="cuda:0"):
    """GPU ReplayMemory holding exactly the oracle replay's arrays (through extend())."""
    from pytorch_r2d2_b200 import ReplayMemory
    g = ReplayMemory(rep.M, batch_size, rep.n_step, cell_size=rep.cell, device=device)
    g.burn_in_length, g.learning_length, g.sequence_length = rep.burn_in, rep.learn, rep.T
    g.extend({k: v.copy() for k, v in rep.memory.items()})
    g.index = rep.index
    return g


def gpu_mem_np(g, key):
    return dict.__getitem__(g.memory, key).cpu().numpy()
