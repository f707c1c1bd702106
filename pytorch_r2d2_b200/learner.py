"""``Learner`` behind the reference's interface (/root/reference learner.py:18-124).

``train()`` is the hot path (learner.py:68-104): PER sample -> burn-in + learn + double-Q
unrolls of the online net, burn-in + evaluation of the target net -> n-step double-Q target,
TD error, loss -> backward -> centered RMSprop -> per-step and per-sequence priority update.
Every stage is a CUDA kernel enqueued through the C ABI on torch's current stream; nothing
crosses to the host inside a step except the B float64 uniforms going in.

Same attribute names and defaults as the reference (``batch_size=8``, 10+10 steps, n=3, ...);
they are plain attributes and can be changed after construction, exactly like tests of the
reference have to do.  Extra, opt-in (defaults reproduce the reference): ``cell_size``,
``rescale`` (invertible value rescaling, absent from the reference), ``world_size``/``rank`` for
the data-parallel gradient allreduce.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import _lib, engine, parallel
from .engine import Weights
from .model import QNet
from .params import FEAT, N_ACTIONS
from .replay_memory import ReplayMemory


def learner_process(n_actors, shared_dict):
    Learner(n_actors, shared_dict).run()


class FusedCenteredRMSprop:
    """torch.optim.RMSprop(lr, alpha, eps, centered=True) (learner.py:51) as ONE kernel over the
    flat parameter buffer."""

    def __init__(self, net: QNet, lr=0.00025 / 4.0, alpha=0.95, eps=1.5e-7):
        self.net, self.lr, self.alpha, self.eps = net, lr, alpha, eps
        flat = net.flat_parameters()
        self.grad = torch.zeros_like(flat)
        self.square_avg = torch.zeros_like(flat)
        self.grad_avg = torch.zeros_like(flat)
        self.grad_scale = 1.0
        self.peer = None
        self._done = set()
        # the same three ranges the data-parallel exchange uses (see attach_peer_exchange): a single
        # replica updates the heads and the LSTM as soon as their gradients are complete, too
        lay, n = net.layout, flat.numel()
        split, head = lay.offset("lstm.weight_ih"), lay.offset("val.0.weight")
        self.buckets = [(0, split), (split, head - split), (head, n - head)]

    def zero_grad(self):
        self.grad.zero_()

    def step(self):
        if self.peer is not None:
            # gradient sum over the ranks + optimiser + parameter all-gather in one kernel per
            # bucket; buckets already exchanged during the backward pass are skipped
            for k in range(len(self.buckets)):
                if k not in self._done:
                    self.step_bucket(k)
            self._done.clear()
            return
        if self._done:                       # some ranges were updated during the backward pass
            for k in range(len(self.buckets)):
                if k not in self._done:
                    self.step_bucket(k)
            self._done.clear()
            return
        engine.rmsprop_centered(self.net.flat_parameters(), self.grad, self.square_avg,
                                self.grad_avg, self.lr, self.alpha, self.eps, self.grad_scale)

    def step_bucket(self, k):
        """Exchange + update bucket k now, on the current stream (its gradients must be complete).
        Single replica: the optimiser kernel on that range of the flat vectors."""
        off, n = self.buckets[k]
        if self.peer is None:
            sl = slice(off, off + n)
            engine.rmsprop_centered(self.net.flat_parameters()[sl], self.grad[sl], self.square_avg[sl],
                                    self.grad_avg[sl], self.lr, self.alpha, self.eps, self.grad_scale)
        else:
            self.peer.reduce_rmsprop(self.square_avg[k], self.grad_avg[k], self.lr, self.alpha, self.eps,
                                     offset=off, n=n, bucket=k)
        self._done.add(k)

    def attach_peer_exchange(self, peer):
        """Switch to the fused NVLink reduce + RMSprop path: parameters and gradients move into the
        IPC-mapped buffers, the optimiser state shrinks to this rank's shard."""
        if self.peer is not None or self.square_avg.abs().max().item() != 0 or self.grad_avg.abs().max().item() != 0:
            raise RuntimeError("attach_peer_exchange must be called once, before the first optimiser step")
        self.net.rebase(peer.param)
        self.grad = peer.grad
        # three buckets in the order the backward pass completes them: [2] the dueling heads (done
        # before the BPTT starts: exchanged beside it), [1] the LSTM (exchanged beside the conv
        # backward), [0] the conv torso (its gradients are the last to complete)
        split = self.net.layout.offset("lstm.weight_ih")
        head = self.net.layout.offset("val.0.weight")
        assert split % 4 == 0 and head % 4 == 0 and split < head < peer.n
        self.buckets = [(0, split), (split, head - split), (head, peer.n - head)]
        z = lambda n: torch.zeros(parallel.shard_elems(n, peer.world), dtype=torch.float32, device=peer.device)  # noqa: E731
        self.square_avg = [z(n) for _, n in self.buckets]
        self.grad_avg = [z(n) for _, n in self.buckets]
        self._done = set()
        self.peer = peer

    def state_dict(self):
        if self.peer is not None:     # sharded: this rank's shards of every bucket
            return {"square_avg": [t.clone() for t in self.square_avg], "grad_avg": [t.clone() for t in self.grad_avg],
                    "lr": self.lr, "alpha": self.alpha, "eps": self.eps, "rank": self.peer.rank, "world": self.peer.world}
        return {"square_avg": self.square_avg.clone(), "grad_avg": self.grad_avg.clone(),
                "lr": self.lr, "alpha": self.alpha, "eps": self.eps}

    def load_state_dict(self, sd):
        if self.peer is not None:
            for dst, src in zip(self.square_avg + self.grad_avg, sd["square_avg"] + sd["grad_avg"]):
                dst.copy_(src)
            self.lr, self.alpha, self.eps = sd["lr"], sd["alpha"], sd["eps"]
            return
        self.square_avg.copy_(sd["square_avg"])
        self.grad_avg.copy_(sd["grad_avg"])
        self.lr, self.alpha, self.eps = sd["lr"], sd["alpha"], sd["eps"]


class Learner:
    def __init__(self, n_actors, shared_dict, device="cuda:0", *, cell_size=256, memory_size=500000,
                 rescale=False, process_group=None, precision="fp32"):
        # params (learner.py:20-28)
        self.gamma = 0.99
        self.alpha = 0.6
        self.bootstrap_steps = 3
        self.initial_exploration = 50000
        self.priority_epsilon = 1e-6
        self.device = device
        self.n_epochs = 0
        self.n_actors = n_actors
        self.rescale = rescale
        self.rescale_eps = 1e-3
        # "fp32": every contraction in fp32 FFMA (1e-5 class vs the reference);
        # "bf16": dense contractions on tcgen05 tensor cores, bf16 operands / fp32 accumulate
        #         (2e-2 class, BASELINE.json north_star)
        if precision not in ("fp32", "bf16"):
            raise ValueError("precision must be 'fp32' or 'bf16'")
        self.precision = precision
        # bf16 mode: run the target net's branch on a side stream next to the online net's
        # (False = one stream, e.g. to time kernels in isolation)
        self.overlap_streams = True
        # run the step's critical chain (online net, data gradients) on a high-priority stream:
        # when its kernels and the side streams' are ready together, its CTAs get the SMs first
        self.prioritize_critical = True
        # capture the whole step (both streams) in a CUDA graph and replay it: removes ~110
        # launches worth of host work per step.  Only the default path (u drawn from the global
        # CPU generator, no overrides) is graphed; anything else runs eagerly.
        self.use_cuda_graph = False
        # also capture the NCCL gradient all-reduce (world > 1).  EXPERIMENTAL and off: capturing it
        # next to the cooperative LSTM launches hung on 2 GPUs; multi-GPU steps run eagerly.
        self.graph_collectives = False
        self._graph, self._graph_key = None, None
        # kernels of this library inside the captured step / replays so far (a replay launches
        # them without passing through the C entry points, so r2d2_launch_count() cannot see them)
        self.graph_kernel_nodes, self.graph_replays = 0, 0
        self.memory_path = os.path.join("./", "logs", "memory")
        # memory (learner.py:34-41)
        self.burn_in_length = 10
        self.learning_length = 10
        self.sequence_length = self.burn_in_length + self.learning_length
        self.memory_size = memory_size
        self.batch_size = 8
        self.memory_load_interval = 20
        self.cell_size = cell_size
        self.replay_memory = ReplayMemory(self.memory_size, self.batch_size, self.bootstrap_steps,
                                          cell_size=cell_size, device=device)
        # net (learner.py:44-51)
        self.shared_dict = shared_dict
        self.net_save_interval = 100
        self.target_update_interval = 1000
        self.net = QNet(self.device, cell_size).to(self.device)
        self.target_net = QNet(self.device, cell_size).to(self.device)
        self.target_net.load_state_dict(self.net.state_dict())
        self.save_model()
        self.process_group = process_group
        parallel.broadcast_flat_params(self.net.flat_parameters(), 0, process_group)
        parallel.broadcast_flat_params(self.target_net.flat_parameters(), 0, process_group)
        self.optim = FusedCenteredRMSprop(self.net)
        self._ws = None
        self.last = None  # tensors of the last step (q, y, delta, loss, ...) for inspection

    # ------------------------------------------------------------------ control loop
    def run(self, max_steps=None):
        """learner.py:53-66: train once the replay holds more than ``initial_exploration``
        transitions, ingest actor dumps until then.  ``max_steps`` (not in the reference, which
        loops forever) returns after that many training steps."""
        while max_steps is None or self.n_epochs < max_steps:
            if self.replay_memory.size > self.initial_exploration:
                self.train()
                self.n_epochs += 1
                if self.n_epochs % 100 == 0:
                    print("trained", self.n_epochs, "epochs")
                self.interval()
            else:
                self._ingest()

    def _ingest(self):
        """Pull every actor's waiting dump into the ring (learner.py:62-66,111-116)."""
        for i in range(self.n_actors):
            try:
                self.replay_memory.load(self.memory_path, i)
            except Exception:  # the reference ignores hand-off errors (learner.py:63-66)
                pass

    def interval(self):
        """learner.py:106-120: target sync every 1000 steps, weight publish every 100, replay
        ingest every 20, weight file every 10000."""
        if self.n_epochs % self.target_update_interval == 0:
            self.sync_target()
        if self.n_epochs % self.net_save_interval == 0:
            self.save_model()
        if self.n_epochs % self.memory_load_interval == 0:
            self._ingest()
        if self.n_epochs % 10000 == 0:
            os.makedirs("save", exist_ok=True)
            torch.save(self.net.cpu_state_dict(), "save/" + str(self.n_epochs) + "_save.pt")

    def sync_target(self):
        """target_net.load_state_dict(net.state_dict()) (learner.py:108) = one D2D copy."""
        self.target_net.flat_parameters().copy_(self.net.flat_parameters())

    def save_model(self):
        """Publish both nets to the actors as CPU state dicts (learner.py:122-124)."""
        self.shared_dict["net_state"] = self.net.cpu_state_dict()
        self.shared_dict["target_net_state"] = self.target_net.cpu_state_dict()

    # ------------------------------------------------------------------ checkpoint / resume
    def _ckpt_path(self, path):
        world = parallel.active_world(self.process_group)
        if world == 1:
            return path
        import torch.distributed as dist
        return f"{path}.rank{dist.get_rank(self.process_group)}"

    def save_checkpoint(self, path, include_replay=True):
        """Everything a resumed run needs to continue bit-identically (the reference only ever
        saves the online weights, learner.py:117-120): both nets, the centered-RMSprop state
        (this rank's shards under the peer exchange), the step counter, the host RNG states the
        sampler draws from, and -- with `include_replay` -- the replay shard (its ten arrays up
        to ``size`` plus the write counter; the sum-tree is rebuilt from them on load).  One file
        per rank (``path.rank<r>``) when data-parallel: every rank owns a replay shard."""
        rm = self.replay_memory
        rm._wait_ingest()
        ck = {"format": 1, "n_epochs": self.n_epochs, "cell_size": self.cell_size,
              "net": self.net.cpu_state_dict(), "target_net": self.target_net.cpu_state_dict(),
              "optim": {k: ([t.cpu() for t in v] if isinstance(v, list) else v.cpu() if torch.is_tensor(v) else v)
                        for k, v in self.optim.state_dict().items()},
              "shape": {"batch_size": self.batch_size, "burn_in_length": self.burn_in_length,
                        "learning_length": self.learning_length, "bootstrap_steps": self.bootstrap_steps},
              "torch_rng": torch.get_rng_state(), "numpy_rng": np.random.get_state()}
        if include_replay:
            ck["replay"] = {"index": rm.index, "memory_size": rm.memory_size, "memory": rm.dump()}
        path = self._ckpt_path(path)
        tmp = path + ".tmp"
        torch.save(ck, tmp)
        os.replace(tmp, path)
        return path

    def load_checkpoint(self, path):
        """Inverse of ``save_checkpoint`` into an already constructed learner of the same width
        (and, under the peer exchange, the same world size)."""
        ck = torch.load(self._ckpt_path(path), weights_only=False)
        if ck.get("format") != 1 or ck["cell_size"] != self.cell_size:
            raise ValueError("checkpoint does not match this learner (format / cell_size)")
        dev = self.net.flat_parameters().device
        self.net.load_state_dict({k: v.to(dev) for k, v in ck["net"].items()})
        self.target_net.load_state_dict({k: v.to(dev) for k, v in ck["target_net"].items()})
        od = ck["optim"]
        if self.optim.peer is not None and (od.get("world") != self.optim.peer.world or od.get("rank") != self.optim.peer.rank):
            raise ValueError("sharded optimiser state was saved by another rank / world size")
        self.optim.load_state_dict({k: ([t.to(dev) for t in v] if isinstance(v, list) else v.to(dev) if torch.is_tensor(v) else v)
                                    for k, v in od.items()})
        self.n_epochs = ck["n_epochs"]
        for k, v in ck["shape"].items():
            setattr(self, k, v)
        self.sequence_length = self.burn_in_length + self.learning_length
        torch.set_rng_state(ck["torch_rng"])
        np.random.set_state(ck["numpy_rng"])
        if "replay" in ck:
            rm, rp = self.replay_memory, ck["replay"]
            if rp["memory_size"] != rm.memory_size:
                raise ValueError("replay shard was saved with another memory_size")
            rm._wait_ingest()
            n = rp["memory"]["state"].shape[0]
            for k, v in rp["memory"].items():
                rm._raw(k)[:n].copy_(torch.from_numpy(v).to(dev).reshape(rm._raw(k)[:n].shape))
            rm.index = rp["index"]
            rm._tree_dirty = True
        self._graph = None

    # ------------------------------------------------------------------ workspace
    def _workspace(self):
        B, BI, L, n, H = (self.batch_size, self.burn_in_length, self.learning_length,
                          self.bootstrap_steps, self.cell_size)
        key = (B, BI, L, n, H, self.precision)
        if self._ws is not None and self._ws["key"] == key:
            return self._ws
        dev = self.net.flat_parameters().device
        T, S = BI + L, BI + L + n
        f = lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)  # noqa: E731
        M = L * B
        ws = dict(key=key)
        ws["a1_on"], ws["a2_on"], ws["feat_on"] = f(S * B, 32, 20, 20), f(S * B, 32, 9, 9), f(S * B, FEAT)
        ws["a1_tg"], ws["a2_tg"], ws["feat_tg"] = f(T * B, 32, 20, 20), f(T * B, 32, 9, 9), f(T * B, FEAT)
        ws["xp_on"], ws["xp_tg"] = f(S * B, 4 * H), f(T * B, 4 * H)
        # one history for both online passes: the double-Q pass continues from the state after
        # state[T-1] (learner.py:86,91), so its slot 0 IS slot T of the first pass -- no copy
        hs_all, cs_all = f(T + 1 + L, B, H), f(T + 1 + L, B, H)
        ws["hs_on"], ws["cs_on"] = hs_all[:T + 1], cs_all[:T + 1]
        ws["hs_dq"], ws["cs_dq"] = hs_all[T:], cs_all[T:]
        ws["hs_tg"], ws["cs_tg"] = f(T + 1, B, H), f(T + 1, B, H)
        ws["gates"] = f(T, B, 4 * H)
        ws["hid_on"], ws["hid_tmp"], ws["hid_tg"] = f(M, 2 * H), f(M, 2 * H), f(M, 2 * H)
        ws["q_all"], ws["qn_on"], ws["qn_tg"] = f(M, N_ACTIONS), f(M, N_ACTIONS), f(M, N_ACTIONS)
        ws["td"] = dict(next_action=torch.empty(M, dtype=torch.int64, device=dev), y=f(M), delta=f(M),
                        dq=f(M, N_ACTIONS), prio=f(M), loss=f(1))
        ws["dhid"], ws["dhs"], ws["dgates"] = f(M, 2 * H), f(M, H), f(M, 4 * H)
        ws["dc_work"], ws["dfeat"] = f(B, H), f(M, FEAT)
        nbytes = max(engine.linear_wgrad_scratch_bytes(M, 2 * H, H),
                     engine.linear_wgrad_scratch_bytes(M, 4 * H, H),
                     engine.linear_wgrad_scratch_bytes(M, 4 * H, FEAT),
                     engine.torso_bwd_scratch_bytes(M))
        ws["scratch"] = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        if self.precision == "bf16":
            if M % 8 != 0:
                raise ValueError("bf16 mode needs learning_length*batch_size to be a multiple of 8 "
                                 "(TMA row pitch of the transposed gradient operands)")
            h = lambda *shape: torch.empty(shape, dtype=torch.bfloat16, device=dev)  # noqa: E731
            # forward operands as bf16 (hi, lo) pairs, stacked on a leading axis of 2
            ws["feat16_on"], ws["feat16_tg"] = h(2, S * B, FEAT), h(2, T * B, FEAT)
            ws["w_ih16"], ws["w_ih16_t"] = h(2, 4 * H, FEAT), h(2, 4 * H, FEAT)
            ws["hw1_16"], ws["hw1_16_t"] = h(2, 2 * H, H), h(2, 2 * H, H)
            ws["hs16"], ws["hs16_t"], ws["hs16_l"] = h(2, M, H), h(2, M, H), h(2, M, H)
            # transposed input projections [4H, frames*B] for the persistent LSTM kernel
            ws["xpT_on"], ws["xpT_tg"] = f(4 * H, S * B), f(4 * H, T * B)
            ws["lstm_scratch"] = torch.empty(engine.tc_lstm_scratch_bytes(B, H), dtype=torch.uint8, device=dev)
            ws["lstm_scratch_tg"] = torch.empty_like(ws["lstm_scratch"])
            ws["lstm_bwd_scratch"] = torch.empty(engine.tc_lstm_bwd_scratch_bytes(B, H), dtype=torch.uint8, device=dev)
            nsm = int(_lib.load().r2d2_sm_count())
            ws["relu1_bits"] = torch.empty((S * B, 400), dtype=torch.int32, device=dev)
            ws["wpart"] = [f(nsm * (32 * 256 + 32)), f(nsm * (32 * 512 + 32)), f(nsm * (32 * 288 + 32))]
            ws["cpart"] = [f(296 * 32), f(296 * 32), f(296 * 32)]   # R2D2_CHANSUM_PART_FLOATS
            ws["side_stream"] = torch.cuda.Stream(device=dev)
            ws["crit_stream"] = torch.cuda.Stream(device=dev, priority=-1)
            ws["aux2_stream"] = torch.cuda.Stream(device=dev)
            ws["xchg_stream"] = torch.cuda.Stream(device=dev, priority=-1)   # multi-GPU exchange: peers wait on it
            ws["aux_stream"] = torch.cuda.Stream(device=dev, priority=-1)   # LSTM weight gradients: short CTAs, gate the exchange
            ws["ev_aux"] = [torch.cuda.Event() for _ in range(3)]
            ws["ev_wprep"] = torch.cuda.Event()
            ws["events"] = [torch.cuda.Event() for _ in range(16)]
            ws["ev_fork"], ws["ev_join"] = torch.cuda.Event(), torch.cuda.Event()
            ws["dhid16"], ws["dg16"], ws["hprev16"] = h(M, 2 * H), h(M, 4 * H), h(M, H)
        self._ws = ws
        return ws

    def enable_peer_exchange(self) -> bool:
        """Data-parallel on one node (one process per GPU): replace the NCCL all-reduce + per-rank
        RMSprop by the fused NVLink reduce + optimiser kernel (csrc/dp_peer.cu).  Call on every
        rank before the first train(); returns False (and changes nothing) for a single process."""
        if parallel.active_world(self.process_group) < 2:
            return False
        try:
            peer = parallel.PeerExchange(self.net.layout.total, self.device, self.process_group)
        except parallel.PeerUnavailable as e:       # raised on every rank together: NCCL path stays
            import warnings
            warnings.warn(f"peer exchange unavailable, keeping the NCCL all-reduce: {e}")
            return False
        self.optim.attach_peer_exchange(peer)
        self._graph = None
        return True

    # ------------------------------------------------------------------ the hot path
    def train(self, u=None, seq_index=None, next_action_override=None):
        """One learner step (learner.py:68-104).  `u` optionally supplies the B float64
        uniforms (default: drawn from torch's global CPU generator like the reference's
        multinomial); `seq_index` bypasses the draw (parity tests)."""
        if (self.use_cuda_graph and u is None and seq_index is None and next_action_override is None
                and (parallel.active_world(self.process_group) == 1 or self.graph_collectives
                     or self.optim.peer is not None)):
            return self._train_graphed()
        return self._train_eager(u, seq_index, next_action_override)

    def _train_graphed(self):
        rm = self.replay_memory
        key = (self.batch_size, self.burn_in_length, self.learning_length, self.bootstrap_steps,
               self.cell_size, self.precision, self.overlap_streams, id(rm))
        rm.configure(self.batch_size, self.bootstrap_steps, self.burn_in_length, self.learning_length)
        # host RNG -> pinned slot -> device, every step, OUTSIDE the graph: a memcpy node would read
        # one fixed host address that the host (running many replays ahead) keeps overwriting
        u_dev = rm._workspace(rm.batch_size, rm.sequence_length, rm.n_step, rm.cell_size)["u"]
        rm.upload_uniforms(self.batch_size, u_dev)
        rm._wait_ingest()                                    # cross-stream waits stay outside the graph
        if self._graph is None or self._graph_key != key:
            # the captured step always rebuilds the sum-tree first (cheap: M*5 bytes), so that
            # replays are valid whether or not rows were ingested / poked since the last step
            rm._tree_dirty = True
            self._train_eager(u_dev, None, None)             # warm-up: allocates every workspace
            torch.cuda.synchronize()
            rm._tree_dirty = True
            g = torch.cuda.CUDAGraph()
            n0 = _lib.launch_count()
            with torch.cuda.graph(g):
                loss = self._train_eager(u_dev, None, None)
            self.graph_kernel_nodes = _lib.launch_count() - n0
            self._graph, self._graph_key, self._graph_loss = g, key, loss
            rm._tree_dirty = False
            return loss      # the capture pass does not execute; the warm-up step above did
        self._graph.replay()
        self.graph_replays += 1
        rm._tree_dirty = False
        return self._graph_loss

    def _train_eager(self, u=None, seq_index=None, next_action_override=None):
        if self.precision == "bf16" and self.overlap_streams and self.prioritize_critical:
            ws = self._workspace()
            outer, crit = torch.cuda.current_stream(), ws["crit_stream"]
            crit.wait_stream(outer)
            with torch.cuda.stream(crit):
                loss = self._train_step(u, seq_index, next_action_override)
            outer.wait_stream(crit)
            return loss
        return self._train_step(u, seq_index, next_action_override)

    def _train_step(self, u=None, seq_index=None, next_action_override=None):
        rm = self.replay_memory
        B, BI, L, n, H = (self.batch_size, self.burn_in_length, self.learning_length,
                          self.bootstrap_steps, self.cell_size)
        T = BI + L
        rm.configure(B, n, BI, L)
        ws = self._workspace()
        bf16 = self.precision == "bf16"
        # tensor-core path: the first conv layer (forward and weight gradient) reads the sampled
        # frames in place from the ring, so the 154 MB slab copy never happens
        w, wt = self.net._weights(), self.target_net._weights()
        if bf16:
            # weight-only operand preparation (bf16 hi/lo splits for the forward GEMMs; the "hi" parts
            # also serve the backward data gradients): nothing here depends on the batch, so it
            # runs on the third stream from the very start of the step
            def weight_prep():
                engine.split_bf16(w.w_ih, ws["w_ih16"][0], ws["w_ih16"][1])
                engine.split_bf16(w.head_w1, ws["hw1_16"][0], ws["hw1_16"][1])
                engine.split_bf16(wt.w_ih, ws["w_ih16_t"][0], ws["w_ih16_t"][1])
                engine.split_bf16(wt.head_w1, ws["hw1_16_t"][0], ws["hw1_16_t"][1])

            if self.overlap_streams:
                ws["ev_aux"][2].record(torch.cuda.current_stream())
                ws["aux_stream"].wait_event(ws["ev_aux"][2])
                with torch.cuda.stream(ws["aux_stream"]):
                    weight_prep()
                    ws["ev_wprep"].record(ws["aux_stream"])
            else:
                weight_prep()
        # the gather writes the initial recurrent states straight into slot 0 of the histories
        batch = rm.sample_slab(u=u, seq_index=seq_index, frames=not bf16,
                               hidden_out=(ws["hs_on"][0], ws["cs_on"][0], ws["hs_tg"][0], ws["cs_tg"][0]))
        slab = batch["slab"]
        g = Weights(self.optim.grad, self.net.layout)
        M = L * B

        bf16 = self.precision == "bf16"
        hs_learn = ws["hs_on"][1 + BI:].reshape(M, H)

        def split(x, pair):
            engine.split_bf16(x, pair[0], pair[1])

        def head(hs, wts, w1_16, hs16, hid, q, presplit=False):
            """Dueling head on [M,H] hidden states (bf16: compensated tcgen05 hidden layer; the
            recurrence kernel has already written the (hi, lo) image when `presplit`)."""
            if bf16:
                if not presplit:
                    split(hs, hs16)
                engine.tc_linear_x3(hs16[0], hs16[1], w1_16[0], w1_16[1], wts.head_b1, None, y32=hid,
                                    relu=True)
                engine.head_out(hid, wts, q)
            else:
                engine.head_fwd(hs, wts, hid, q)

        def target_branch():
            """Target net (learner.py:72,79,92): torso on frames n..T+n, burn-in on
            next_state[0:BI], evaluation on next_state[BI:T].  Independent of the online net
            until the TD target, so in bf16 mode it runs on the side stream."""
            if fused_conv1:
                engine.torso_fwd_tail(ws["a1_tg"], wt, ws["a2_tg"], ws["feat_tg"], ws["feat16_tg"])
            else:
                engine.torso_fwd_u8(slab, n, T, wt, ws["a1_tg"], ws["a2_tg"], ws["feat_tg"], tc=bf16)
            if bf16:
                if overlap:
                    torch.cuda.current_stream().wait_event(ws["ev_wprep"])
                if not fused_conv1:
                    split(ws["feat_tg"], ws["feat16_tg"])
                # the projection kernel is persistent: it leaves the SMs the online net's recurrence
                # (a cooperative launch that is ready while this one runs) needs to start
                free = (H // 16) * ((B + 31) // 32) if overlap else 0
                sms = ws.setdefault("sm_count", int(_lib.load().r2d2_sm_count()))
                engine.tc_linear_x3(ws["w_ih16_t"][0], ws["w_ih16_t"][1], ws["feat16_tg"][0],
                                    ws["feat16_tg"][1], wt.b_ih, wt.b_hh, y32=ws["xpT_tg"], bias_row=True,
                                    max_ctas=sms - free if 0 < free < sms - 32 else 0)
                engine.tc_lstm_seq_fwd(ws["xpT_tg"], 0, wt.w_hh, ws["hs_tg"], ws["cs_tg"], None, T, B, H,
                                       ws["lstm_scratch_tg"], h16=ws["hs16_t"], hist_t0=BI)
            else:
                engine.linear_fwd(ws["feat_tg"], wt.w_ih, wt.b_ih, wt.b_hh, ws["xp_tg"])
                engine.lstm_seq_fwd(ws["xp_tg"], wt.w_hh, ws["hs_tg"], ws["cs_tg"], None, T, B, H)
            head(ws["hs_tg"][1 + BI:].reshape(M, H), wt, ws.get("hw1_16_t"), ws.get("hs16_t"),
                 ws["hid_tg"], ws["qn_tg"], presplit=bf16)

        overlap = bf16 and self.overlap_streams
        # both nets read the same frames and the first layer is bound by gathering them: one pass
        # builds each receptive field once and multiplies it against both nets' filters
        fused_conv1 = bf16
        if fused_conv1:
            engine.conv1_fwd_dual_ring(batch["ring"], batch["seq_index"], T + n, w, ws["a1_on"], wt, ws["a1_tg"], n,
                                       relu_bits_a=ws["relu1_bits"])
        if overlap:
            main, side = torch.cuda.current_stream(), ws["side_stream"]
            ws["ev_fork"].record(main)
            side.wait_event(ws["ev_fork"])
            with torch.cuda.stream(side):
                target_branch()
                ws["ev_join"].record(side)
        else:
            target_branch()

        # ---- online net: torso on all T+n frames; burn-in + learn over state[0:T] (the hidden
        #      state carries through); then again on next_state[BI:T] = frames BI+n.., starting
        #      from the state left after state[T-1] (the reference never resets it:
        #      learner.py:86,91)
        if fused_conv1:
            engine.torso_fwd_tail(ws["a1_on"], w, ws["a2_on"], ws["feat_on"], ws["feat16_on"])
        else:
            engine.torso_fwd_u8(slab, 0, T + n, w, ws["a1_on"], ws["a2_on"], ws["feat_on"], tc=bf16)
        if bf16:
            # forward contractions: error-compensated bf16x3;
            # xproj^T[4H, frames*B] = W_ih . feat^T + (b_ih + b_hh) per row
            if overlap:
                main.wait_event(ws["ev_wprep"])
            if not fused_conv1:
                split(ws["feat_on"], ws["feat16_on"])
            engine.tc_linear_x3(ws["w_ih16"][0], ws["w_ih16"][1], ws["feat16_on"][0], ws["feat16_on"][1],
                                w.b_ih, w.b_hh, y32=ws["xpT_on"], bias_row=True)
            sc = ws["lstm_scratch"]
            engine.tc_lstm_seq_fwd(ws["xpT_on"], 0, w.w_hh, ws["hs_on"], ws["cs_on"], ws["gates"], T, B, H, sc,
                                   h16=ws["hs16_l"], hist_t0=BI)
            if overlap:
                # the learn-window head only needs hs_on: it runs on a third stream beside the
                # double-Q pass of the recurrence (which occupies 64 of the 148 SMs)
                ws["ev_aux"][0].record(main)
                ws["aux_stream"].wait_event(ws["ev_aux"][0])
                with torch.cuda.stream(ws["aux_stream"]):
                    head(hs_learn, w, ws["hw1_16"], ws["hs16_l"], ws["hid_on"], ws["q_all"], presplit=True)
                    ws["ev_aux"][1].record(ws["aux_stream"])
            engine.tc_lstm_seq_fwd(ws["xpT_on"], (BI + n) * B, w.w_hh, ws["hs_dq"], ws["cs_dq"], None, L, B, H, sc,
                                   h16=ws["hs16"], hist_t0=0)
        else:
            engine.linear_fwd(ws["feat_on"], w.w_ih, w.b_ih, w.b_hh, ws["xp_on"])
            engine.lstm_seq_fwd(ws["xp_on"], w.w_hh, ws["hs_on"], ws["cs_on"], ws["gates"], T, B, H)
            engine.lstm_seq_fwd(ws["xp_on"][(BI + n) * B:], w.w_hh, ws["hs_dq"], ws["cs_dq"], None, L, B, H)

        # ---- dueling heads on the learn window only
        head(ws["hs_dq"][1:].reshape(M, H), w, ws.get("hw1_16"), ws.get("hs16"), ws["hid_tmp"], ws["qn_on"],
             presplit=bf16)
        if overlap:
            main.wait_event(ws["ev_aux"][1])
            main.wait_event(ws["ev_join"])
        else:
            head(hs_learn, w, ws.get("hw1_16"), ws.get("hs16_l"), ws["hid_on"], ws["q_all"], presplit=bf16)

        # ---- n-step double-Q target, TD error, loss gradient, per-step priorities
        td = ws["td"]
        engine.td_target(ws["q_all"], ws["qn_on"], ws["qn_tg"], batch["action"].reshape(M),
                         batch["reward"].reshape(M), batch["done"].reshape(M),
                         self.gamma ** self.bootstrap_steps, td, rescale=self.rescale,
                         rescale_eps=self.rescale_eps, prio_eps=self.priority_epsilon,
                         alpha=self.alpha, next_action_override=next_action_override)

        # ---- backward (learn window only; BPTT truncated at the burn-in boundary)
        scratch = ws["scratch"]
        hprev = ws["hs_on"][BI:BI + L].reshape(M, H)
        feat_l = ws["feat_on"][BI * B: T * B]
        a1_l, a2_l = ws["a1_on"][BI * B: T * B], ws["a2_on"][BI * B: T * B]
        if bf16:
            # The data-gradient chain (head -> BPTT -> input projection -> conv) is the critical
            # path; everything that only CONSUMES a gradient (weight/bias gradients, the priority
            # update) runs beside it on the side stream.
            ev = iter(ws["events"])
            main, side = torch.cuda.current_stream(), ws["side_stream"]
            overlap_bwd = self.overlap_streams

            aux = ws["aux_stream"]

            def beside(fn, stream=None):
                """Run fn() on a side stream after everything enqueued so far on main."""
                if not overlap_bwd:
                    return fn()
                stream = stream or side
                e = next(ev)
                e.record(main)
                stream.wait_event(e)
                with torch.cuda.stream(stream):
                    fn()

            def priorities():
                rm.update_priority_learn_window(batch["seq_index"], td["prio"])
                rm.update_sequence_priority(batch["seq_index"], True)

            beside(priorities)
            engine.head_bwd(td["dq"], ws["hid_on"], w, ws["dhid"], None, None)

            engine.cast_bf16(ws["dhid"], ws["dhid16"])

            # weight gradients dW = dY^T . X straight from the row-major bf16 gradients and
            # activations the step already holds (both operands MN-major: no transposed copies)
            def head_wgrad():     # dW2, db2; dW1 = dhid^T . hs
                engine.head_bwd(td["dq"], ws["hid_on"], w, None, g.head_w2, g.head_b2)
                engine.tc_linear_tn(ws["dhid16"], ws["hs16_l"][0], g.head_w1)
                engine.colsum(ws["dhid"], g.head_b1)

            beside(head_wgrad)
            # early optimiser ranges: with the peer exchange, or as a single replica (an NCCL
            # all-reduce of the whole gradient, the fallback, has to wait for the end of the step)
            early = overlap_bwd and (self.optim.peer is not None or parallel.active_world(self.process_group) == 1)
            if early:
                # the heads' gradients are complete: exchanged / applied beside the BPTT (most SMs idle)
                xs = ws["xchg_stream"]
                e = next(ev)
                e.record(side)
                xs.wait_event(e)
                with torch.cuda.stream(xs):
                    self.optim.step_bucket(2)
            # data gradients multiply by the weights as stored: the bf16 "hi" copies of the forward pass
            engine.tc_linear_nn(ws["dhid16"], ws["hw1_16"][0], ws["dhs"])

            def lstm_wgrad_operands():   # forward-only operand: cast while the BPTT runs
                engine.cast_bf16(hprev, ws["hprev16"])

            beside(lstm_wgrad_operands, aux)
            engine.tc_lstm_seq_bwd(ws["dhs"], w.w_hh, ws["gates"][BI:], ws["cs_on"][BI:], ws["dgates"],
                                   L, B, H, ws["lstm_bwd_scratch"], dgates16=ws["dg16"])

            def lstm_wgrad():
                engine.tc_linear_tn(ws["dg16"], ws["hprev16"], g.w_hh)
                engine.tc_linear_tn(ws["dg16"], ws["feat16_on"][0][BI * B: T * B], g.w_ih)
                engine.colsum(ws["dgates"], g.b_hh, g.b_ih)

            beside(lstm_wgrad, aux)       # the longest consumer chain gets a stream of its own
            if early:
                # the LSTM gradients are complete once lstm_wgrad (aux) is: their exchange + update
                # crosses NVLink (or, single replica, just runs) beside the conv backward
                xs = ws["xchg_stream"]
                e = next(ev)
                e.record(aux)
                xs.wait_event(e)
                with torch.cuda.stream(xs):
                    self.optim.step_bucket(1)
            engine.tc_linear_nn(ws["dg16"], ws["w_ih16"][0], ws["dfeat"])
            # conv torso: scratch = [dz2 | dz1 | ...]; dz3 = dfeat masked in place
            sc32 = scratch.view(torch.float32)
            dz2, dz1 = sc32[:M * 2592], sc32[M * 2592: M * (2592 + 12800)]
            dz3 = ws["dfeat"]
            P, st = _lib.ptr, _lib.stream_ptr
            _lib.call("r2d2_relu_mask", P(dz3, "f32"), P(feat_l, "f32"), dz3.numel(), st())

            def conv3_grads():
                # weight AND bias gradient (the bias sums fall out of the operand build)
                _lib.call("r2d2_tc_conv_wgrad_layer", 3, P(a2_l, "f32"), B, 0, 0, L, P(dz3, "f32"), P(g.w3),
                          P(g.b3), P(ws["wpart"][2], "f32"), st())

            beside(conv3_grads)
            _lib.call("r2d2_tc_conv_dgrad3", P(dz3, "f32"), P(w.w3), P(a2_l, "f32"), P(dz2, "f32"), M, st())

            def conv2_grads():
                _lib.call("r2d2_tc_conv_wgrad_layer", 2, P(a1_l, "f32"), B, 0, 0, L, P(dz2, "f32"), P(g.w2),
                          P(g.b2), P(ws["wpart"][1], "f32"), st())

            beside(conv2_grads, ws["aux2_stream"])
            # ReLU' of the first layer from the bit mask its forward epilogue wrote (learn window)
            _lib.call("r2d2_tc_conv_dgrad2_masked", P(dz2, "f32"), P(w.w2),
                      P(ws["relu1_bits"][BI * B:], "i32"), P(dz1, "f32"), M, st())
            engine.conv1_wgrad_ring(batch["ring"], batch["seq_index"], BI, L, dz1, g.w1, ws["wpart"][0], db1=g.b1)
            if overlap_bwd:
                for st_ in (side, aux, ws["aux2_stream"]) + ((ws["xchg_stream"],) if early else ()):
                    e = next(ev)
                    e.record(st_)
                    main.wait_event(e)
        else:
            engine.head_bwd(td["dq"], ws["hid_on"], w, ws["dhid"], g.head_w2, g.head_b2)
            engine.linear_wgrad(ws["dhid"], hs_learn, g.head_w1, g.head_b1, scratch)
            engine.linear_dgrad(ws["dhid"], w.head_w1, ws["dhs"])
            engine.lstm_seq_bwd(ws["dhs"], w.w_hh, ws["gates"][BI:], ws["cs_on"][BI:], ws["dgates"],
                                ws["dc_work"], L, B, H)
            engine.linear_wgrad(ws["dgates"], hprev, g.w_hh, g.b_hh, scratch)
            g.b_ih.copy_(g.b_hh)
            engine.linear_wgrad(ws["dgates"], feat_l, g.w_ih, None, scratch)
            engine.linear_dgrad(ws["dgates"], w.w_ih, ws["dfeat"])
            engine.torso_bwd_u8(slab, BI, L, a1_l, a2_l, feat_l, ws["dfeat"], w, g, scratch)

        # ---- optimiser (data-parallel: sum gradients over ranks, 1/world folded into K10)
        if self.optim.peer is None:
            self.optim.grad_scale = parallel.allreduce_flat_grad(self.optim.grad, self.process_group)
        self.optim.step()

        # ---- priorities (learner.py:102-104); the bf16 path already ran them beside the backward
        if not bf16:
            rm.update_priority_learn_window(batch["seq_index"], td["prio"])
            rm.update_sequence_priority(batch["seq_index"], True)
        self.last = dict(loss=td["loss"], q_all=ws["q_all"], q_next_online=ws["qn_on"],
                         q_next_target=ws["qn_tg"], next_action=td["next_action"], y=td["y"],
                         delta=td["delta"], priority=td["prio"], seq_index=batch["seq_index"],
                         index=batch["index"])
        return td["loss"]
