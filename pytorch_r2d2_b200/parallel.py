"""Data-parallel plumbing of the learner (host logic; works with any torch.distributed backend).

The path shards per sequence (learner.py:98 is a plain mean over L*B elements, no importance
weights; sampling is with replacement, replay_memory.py:230), so each rank owns a replay shard,
samples its own B sequences and the ONLY exchange is the sum of the flat f32 gradient buffer.
Equal per-rank batch sizes make the mean of the shard means equal to the global-batch mean, so
the rank-summed gradient times 1/world is the gradient of the reference's loss at batch world*B.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def world_info():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_from_env(backend: str | None = None):
    """Create the default process group from torchrun's environment (no-op for world 1)."""
    rank, local, world = world_info()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device(f"cuda:{local}")
        dist.init_process_group(backend, **kw)
    return rank, local, world


def active_world(group=None) -> int:
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group)
    return 1


def allreduce_flat_grad(flat_grad: torch.Tensor, group=None) -> float:
    """Sum the flat gradient over ranks in place (one collective per step) and return the
    factor the optimiser kernel must apply (1/world) -- the scale is fused into K10 instead of
    costing another pass over the buffer."""
    world = active_world(group)
    if world == 1:
        return 1.0
    dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    return 1.0 / world


def broadcast_flat_params(flat: torch.Tensor, src: int = 0, group=None) -> None:
    """Replicas must start from identical weights (the reference has one learner, so this is the
    DP analogue of `target_net.load_state_dict(net.state_dict())`, learner.py:49)."""
    if active_world(group) > 1:
        dist.broadcast(flat, src=src, group=group)


def shard_elems(n: int, world: int) -> int:
    """Optimiser-state floats rank r keeps for its shard of an n-element parameter vector
    (mirror of r2d2_dp_shard_elems: float4-granular shards + room for the scalar tail)."""
    per = (n // 4 + world - 1) // world
    return per * 4 + 4


class _DevBuf:
    """A raw device allocation exposed to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, n: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 2}


class PeerUnavailable(RuntimeError):
    """CUDA IPC / peer access could not be set up on every rank (raised on ALL ranks together)."""


class PeerExchange:
    """Gradient / parameter / flag buffers of every rank mapped into this process over CUDA IPC,
    for the fused NVLink reduce + RMSprop kernel (csrc/dp_peer.cu).  One process per GPU on one
    node; the handles travel through torch.distributed (any backend)."""

    def __init__(self, n: int, device, group=None):
        import ctypes
        from . import _lib
        self.n, self.group = n, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        sizes = {"grad": n * 4, "param": n * 4, "flags": 4096}
        own, handles = {}, {}
        try:
            for name, nbytes in sizes.items():
                p, h = ctypes.c_void_p(), ctypes.create_string_buffer(64)
                _lib.call("r2d2_peer_alloc", nbytes, ctypes.byref(p), h)
                own[name], handles[name] = p.value, bytes(h.raw)
            torch.cuda.synchronize()
            ok = True
        except Exception as e:          # noqa: BLE001 -- the decision below must be collective
            ok, why = False, repr(e)
        # every rank must take the same branch: agree before anybody waits for anybody
        if not self._all_ok(ok):
            raise PeerUnavailable(f"rank {self.rank}: IPC export failed on some rank" + ("" if ok else f" ({why})"))
        gathered = [None] * self.world
        dist.all_gather_object(gathered, handles, group=group)
        self._own, self._opened = own, []
        self.ptrs = {name: [0] * self.world for name in sizes}
        try:
            for r in range(self.world):
                for name in sizes:
                    if r == self.rank:
                        self.ptrs[name][r] = own[name]
                    else:
                        p = ctypes.c_void_p()
                        _lib.call("r2d2_peer_open", ctypes.create_string_buffer(gathered[r][name], 64), ctypes.byref(p))
                        self.ptrs[name][r] = p.value
                        self._opened.append(p.value)
            ok = True
        except Exception as e:          # noqa: BLE001
            ok, why = False, repr(e)
        if not self._all_ok(ok):
            raise PeerUnavailable(f"rank {self.rank}: mapping a peer's buffers failed on some rank (no P2P / IPC "
                                  "between these devices?)" + ("" if ok else f" ({why})"))
        arr = ctypes.c_void_p * self.world
        self.c_ptrs = {name: arr(*self.ptrs[name]) for name in sizes}
        self.grad = torch.as_tensor(_DevBuf(own["grad"], n, "<f4"), device=self.device)
        self.param = torch.as_tensor(_DevBuf(own["param"], n, "<f4"), device=self.device)
        dist.barrier(group=group)

    def _all_ok(self, ok: bool) -> bool:
        t = torch.tensor([1 if ok else 0], dtype=torch.int32,
                         device=self.device if dist.get_backend(self.group) == "nccl" else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return bool(t.item())

    def reduce_rmsprop(self, square_avg_shard, grad_avg_shard, lr, alpha, eps, offset=0, n=None, bucket=0):
        """Exchange + optimiser step for the slice [offset, offset+n) of the flat vectors (default:
        all of it); every slice needs its own `bucket` id and its own state shards."""
        from . import _lib
        n = self.n - offset if n is None else n
        _lib.call("r2d2_dp_reduce_rmsprop", self.rank, self.world, self.c_ptrs["grad"], self.c_ptrs["param"],
                  self.c_ptrs["flags"], _lib.ptr(square_avg_shard, "f32"), _lib.ptr(grad_avg_shard, "f32"),
                  offset, bucket, n, float(lr), float(alpha), float(eps), 1.0 / self.world, _lib.stream_ptr())

    def close(self):
        from . import _lib
        torch.cuda.synchronize()
        for p in self._opened:
            _lib.call("r2d2_peer_close", p)
        self._opened = []
