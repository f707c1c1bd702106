"""CPU, world_size 2, gloo: the data-parallel contract of the learner.
Each rank computes the oracle gradient of ITS half of the batch; the rank-summed flat gradient
times the factor returned by allreduce_flat_grad must equal the oracle gradient of the whole
batch (mean of shard means == global mean), and replicas stay bit-identical after the update."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.set_num_threads(2)
    import oracle
    oracle.warm_up()                    # spawned process: see oracle.warm_up
    from oracle import net_oracle as NO
    from pytorch_r2d2_b200 import parallel
    from pytorch_r2d2_b200.params import MODULE_ORDER, ParamLayout
    from util import make_oracle_replay
    r, _, w = parallel.init_from_env("gloo")
    assert (r, w) == (rank, world)
    H, B, BI, LN, NS = 256, 2, 2, 2, 1
    rep, _ = make_oracle_replay(48, H, NS, BI, LN, seed=5)
    p = NO.init_params(0, H)
    lay = ParamLayout(H)
    flat = torch.zeros(lay.total)
    if rank == 0:
        for k in MODULE_ORDER:
            lay.view(flat, k).copy_(p[k])
    parallel.broadcast_flat_params(flat, 0)                      # replicas start identical
    p = {k: lay.view(flat, k).clone() for k in MODULE_ORDER}
    seq_all = np.array([0, 9, 21, 30])                           # global batch of 2*B sequences
    mine = seq_all[rank * B:(rank + 1) * B]
    res = NO.train_step(p, p, NO.to_torch_batch(rep.gather(mine)["batch"]), burn_in=BI, n_step=NS)
    g = torch.zeros(lay.total)
    for k in MODULE_ORDER:
        lay.view(g, k).copy_(res["grads"][k])
    scale = parallel.allreduce_flat_grad(g)
    assert scale == 0.5
    g *= scale
    full = NO.train_step(p, p, NO.to_torch_batch(rep.gather(seq_all)["batch"]), burn_in=BI, n_step=NS)
    for k in MODULE_ORDER:
        want = full["grads"][k]
        err = float((lay.view(g, k) - want).abs().max() / want.abs().max().clamp_min(1e-30))
        assert err < 1e-5, (k, err)
    sq, ga = torch.zeros_like(flat), torch.zeros_like(flat)
    NO.rmsprop_centered(flat, g, sq, ga)
    torch.save(flat, os.path.join(out_dir, f"flat{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_dp_gradient_equals_global_batch(tmp_path):
    world, port = 2, 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    a, b = torch.load(tmp_path / "flat0.pt"), torch.load(tmp_path / "flat1.pt")
    assert torch.equal(a, b)


def test_single_process_is_identity():
    sys.path.insert(0, ROOT)
    from pytorch_r2d2_b200 import parallel
    g = torch.arange(8.0)
    assert parallel.allreduce_flat_grad(g) == 1.0 and torch.equal(g, torch.arange(8.0))
    assert parallel.active_world() == 1
