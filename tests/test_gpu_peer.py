"""Fused NVLink gradient reduce + RMSprop (csrc/dp_peer.cu) on two GPUs: needs >= 2 devices, so it
is skipped on a single-GPU box (run it with `gpurun --gpus 2 -- python -m pytest tests/test_gpu_peer.py -m gpu`)."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pytorch_r2d2_b200 import Learner, parallel
    dev = f"cuda:{rank}"
    # ---- kernel level: three steps against a local torch restatement (every rank can rebuild
    #      every rank's gradient from its seed)
    n, lr, alpha, eps = 100_003, 1e-3, 0.95, 1.5e-7
    split = 40_000                      # two buckets, exchanged separately (bucket 1 first)
    pe = parallel.PeerExchange(n, dev)
    g0 = torch.Generator().manual_seed(7)
    p_ref = torch.randn(n, generator=g0).double()
    pe.param.copy_(p_ref.float())
    v = [torch.zeros(parallel.shard_elems(k, world), device=dev) for k in (split, n - split)]
    m = [torch.zeros_like(x) for x in v]
    v_ref, m_ref = torch.zeros(n).double(), torch.zeros(n).double()
    p_ref = p_ref.float().double()
    dist.barrier()
    for step in range(3):
        grads = [torch.randn(n, generator=torch.Generator().manual_seed(100 * step + r)) for r in range(world)]
        pe.grad.copy_(grads[rank])
        pe.reduce_rmsprop(v[1], m[1], lr, alpha, eps, offset=split, n=n - split, bucket=1)
        pe.reduce_rmsprop(v[0], m[0], lr, alpha, eps, offset=0, n=split, bucket=0)
        torch.cuda.synchronize()
        g = sum(x.double() for x in grads) / world
        v_ref = alpha * v_ref + (1 - alpha) * g * g
        m_ref = m_ref + (1 - alpha) * (g - m_ref)
        p_ref = p_ref - lr * g / ((v_ref - m_ref * m_ref).sqrt() + eps)
        err = (pe.param.cpu().double() - p_ref).abs().max().item() / p_ref.abs().max().item()
        assert err < 1e-5, (step, err)
    mine = pe.param.cpu()
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    assert all(torch.equal(gathered[0], x) for x in gathered), "replicas diverged"
    pe.close()
    # ---- learner level: two data-parallel steps, replicas stay bit-identical
    from util import make_oracle_replay, mirror_to_gpu
    torch.manual_seed(0)
    L = Learner(1, {}, device=dev, cell_size=256, memory_size=64, precision="bf16")
    L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = 8, 3, 4, 2
    rep, _ = make_oracle_replay(64, 256, 2, 3, 4, seed=21 + rank)
    L.replay_memory = mirror_to_gpu(rep, 8, device=dev)
    assert L.enable_peer_exchange()
    losses = [float(L.train()) for _ in range(2)]
    torch.cuda.synchronize()
    assert all(x == x for x in losses)
    mine = L.net.flat_parameters().cpu()
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    assert all(torch.equal(gathered[0], x) for x in gathered), "learner replicas diverged"
    out[rank] = losses
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_reduce_rmsprop_two_gpus():
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, 29533, out), nprocs=2, join=True)
    assert len(out) == 2
