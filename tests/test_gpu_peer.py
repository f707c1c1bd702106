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
    # ---- learner level.  Step 1 against the oracle: every rank runs the oracle's train step on its
    #      OWN 8 sequences; with equal shard batches the mean of the per-rank oracle gradients is
    #      the reference's gradient at the 16-sequence global batch.  The
    #      CUDA side of that comparison: the local gradients the exchange kernel reduces (still in
    #      the IPC buffers after the step), averaged over the ranks -- within the bf16 class -- and
    #      the parameters every replica ends up with == centered RMSprop of that average.
    import numpy as np
    import oracle
    oracle.warm_up()                    # spawned process: see oracle.warm_up
    from oracle import net_oracle as NO
    from util import make_oracle_replay, mirror_to_gpu, rel_err
    torch.manual_seed(0)
    B, BI, LN, NS, H = 8, 3, 4, 2, 256
    L = Learner(1, {}, device=dev, cell_size=H, memory_size=64, precision="bf16")
    L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = B, BI, LN, NS
    rep, _ = make_oracle_replay(64, H, NS, BI, LN, seed=21 + rank)
    L.replay_memory = mirror_to_gpu(rep, B, device=dev)
    assert L.enable_peer_exchange()
    p0 = {k: v.cpu().clone() for k, v in L.net.state_dict().items()}
    pt0 = {k: v.cpu().clone() for k, v in L.target_net.state_dict().items()}
    u = np.random.RandomState(50 + rank).rand(B)
    si = rep.sample_indices(u)
    res = NO.train_step(p0, pt0, NO.to_torch_batch(rep.gather(si)["batch"]), burn_in=BI, n_step=NS)
    flat0 = L.net.flat_parameters().cpu().clone()
    loss1 = float(L.train(u=torch.from_numpy(u), next_action_override=res["next_action"].to(dev)))
    torch.cuda.synchronize()
    assert np.array_equal(L.last["seq_index"].cpu().numpy(), si)
    lay = L.net.layout
    local = L.optim.grad.cpu().clone()                       # this rank's gradient, as exchanged
    # 32 learn-window frames: one ReLU decision on the other side of a kink moves a conv gradient
    # by ~1e-2, so the oracle gradient is taken with the GPU's activation pattern pinned
    # (tests/test_gpu_fullshape.py explains and bounds the flips)
    ws, T = L._ws, BI + LN
    masks = [(ws[k][BI * B:T * B] > 0).float().cpu() for k in ("a1_on", "a2_on", "feat_on")]
    masks[2] = masks[2].reshape(-1, 32, 7, 7)
    tb = NO.to_torch_batch(rep.gather(si)["batch"])
    zs = NO.torso_preacts(p0, tb["state"][BI:].reshape(-1, 4, 84, 84))
    flips = sum(int(((m > 0) != (z > 0)).sum()) for m, z in zip(masks, zs))
    assert flips < 20, flips
    res = NO.train_step(p0, pt0, tb, burn_in=BI, n_step=NS, next_action_override=res["next_action"], learn_masks=masks)
    errs = {k: rel_err(lay.view(local, k), res["grads"][k]) for k in NO.PARAM_KEYS}
    assert max(errs.values()) < 2e-2, (rank, {k: f"{v:.1e}" for k, v in errs.items()})
    both = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(both, local)
    reduced = sum(both) / world                              # what the kernel's rank-ordered sum * 1/world is
    oracle_flat = torch.zeros_like(local)
    for k in NO.PARAM_KEYS:
        lay.view(oracle_flat, k).copy_(res["grads"][k])
    oracles = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(oracles, oracle_flat)
    global_oracle = sum(oracles) / world                     # oracle gradient of the 16-sequence batch
    for k in NO.PARAM_KEYS:
        assert rel_err(lay.view(reduced, k), lay.view(global_oracle, k)) < 2e-2, (rank, k)
    want = flat0.clone()
    NO.rmsprop_centered(want, reduced.clone(), torch.zeros_like(want), torch.zeros_like(want))
    assert rel_err(L.net.flat_parameters().cpu(), want) < 1e-5
    # ---- then free-running: replicas stay bit-identical
    losses = [loss1] + [float(L.train()) for _ in range(2)]
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
