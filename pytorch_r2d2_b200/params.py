"""Flat parameter layout of QNet (host logic, no CUDA needed).

All 18 parameter tensors of the reference network (model.py:12-36) live in ONE contiguous f32
buffer so that the optimiser (K10) and the data-parallel gradient allreduce touch a single
allocation.  The order groups tensors the kernels read together:
``val.0.weight`` / ``adv.0.weight`` are adjacent (one [2H,H] GEMM for both hidden layers of the
dueling head), likewise their biases and the two output layers ([7,H] and [7]).
"""
from __future__ import annotations

from collections import OrderedDict

FEAT = 7 * 7 * 32
N_ACTIONS = 6

# (state_dict key, shape as a function of H) in FLAT order
_FLAT_ORDER = (
    ("vis_layers.0.weight", lambda H: (32, 4, 8, 8)),
    ("vis_layers.0.bias", lambda H: (32,)),
    ("vis_layers.2.weight", lambda H: (32, 32, 4, 4)),
    ("vis_layers.2.bias", lambda H: (32,)),
    ("vis_layers.4.weight", lambda H: (32, 32, 3, 3)),
    ("vis_layers.4.bias", lambda H: (32,)),
    ("lstm.weight_ih", lambda H: (4 * H, FEAT)),
    ("lstm.weight_hh", lambda H: (4 * H, H)),
    ("lstm.bias_ih", lambda H: (4 * H,)),
    ("lstm.bias_hh", lambda H: (4 * H,)),
    ("val.0.weight", lambda H: (H, H)),
    ("adv.0.weight", lambda H: (H, H)),
    ("val.0.bias", lambda H: (H,)),
    ("adv.0.bias", lambda H: (H,)),
    ("val.2.weight", lambda H: (1, H)),
    ("adv.2.weight", lambda H: (N_ACTIONS, H)),
    ("val.2.bias", lambda H: (1,)),
    ("adv.2.bias", lambda H: (N_ACTIONS,)),
)

# state_dict / parameters() order of the reference module (model.py:12-36)
MODULE_ORDER = (
    "vis_layers.0.weight", "vis_layers.0.bias", "vis_layers.2.weight", "vis_layers.2.bias",
    "vis_layers.4.weight", "vis_layers.4.bias",
    "lstm.weight_ih", "lstm.weight_hh", "lstm.bias_ih", "lstm.bias_hh",
    "val.0.weight", "val.0.bias", "val.2.weight", "val.2.bias",
    "adv.0.weight", "adv.0.bias", "adv.2.weight", "adv.2.bias",
)


class ParamLayout:
    """Offsets (in f32 elements) of every tensor inside the flat buffer."""

    def __init__(self, cell: int = 256):
        if cell % 32 != 0:
            raise ValueError("cell size must be a multiple of 32")
        self.cell = cell
        self.entries = OrderedDict()
        off = 0
        for key, shp in _FLAT_ORDER:
            shape = shp(cell)
            n = 1
            for s in shape:
                n *= s
            self.entries[key] = (off, n, shape)
            off += n
        self.total = off

    def offset(self, key):
        return self.entries[key][0]

    def numel(self, key):
        return self.entries[key][1]

    def shape(self, key):
        return self.entries[key][2]

    def view(self, flat, key):
        off, n, shape = self.entries[key]
        return flat[off: off + n].view(shape)

    def span(self, first, last):
        """[offset, numel) covering the adjacent tensors first..last (inclusive)."""
        keys = list(self.entries)
        i, j = keys.index(first), keys.index(last)
        off = self.entries[first][0]
        end = self.entries[last][0] + self.entries[last][1]
        assert i <= j
        return off, end - off


def shard_bounds(total: int, world: int, rank: int):
    """Contiguous, 4-element aligned shard of a flat buffer for rank `rank` of `world`."""
    per = -(-total // world)
    per = -(-per // 4) * 4
    lo = min(rank * per, total)
    hi = min(lo + per, total)
    return lo, hi
