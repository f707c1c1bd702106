"""pytorch_r2d2_b200 -- the learner hot path of LiXirong/pytorch-r2d2 as hand-written sm_100a
CUDA kernels behind the reference's Python API (ReplayMemory, QNet, Learner).

Importing the package does not need a GPU; constructing any of its classes does, and there is
no CPU or PyTorch-eager fallback: a missing ``libr2d2_b200.so`` raises.
"""
from . import _lib  # noqa: F401
from .params import ParamLayout  # noqa: F401


def __getattr__(name):
    if name == "ReplayMemory":
        from .replay_memory import ReplayMemory
        return ReplayMemory
    if name == "QNet":
        from .model import QNet
        return QNet
    if name in ("Learner", "learner_process"):
        from . import learner
        return getattr(learner, name)
    if name in ("BatchedActor", "ActorExperience", "Actor", "actor_process"):
        from . import actor
        return getattr(actor, name)
    raise AttributeError(name)
