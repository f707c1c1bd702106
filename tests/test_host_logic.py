"""CPU: host-side logic and the C-ABI surface (no compute calls)."""
import ctypes
import os
import re

import numpy as np
import pytest

from pytorch_r2d2_b200 import _lib
from pytorch_r2d2_b200.params import MODULE_ORDER, ParamLayout, shard_bounds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "r2d2_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(r2d2_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = header_functions()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/r2d2_b200.h but not exported"
    assert set(names) == set(_lib.PROTOTYPES), set(names) ^ set(_lib.PROTOTYPES)
    assert lib.r2d2_version() >= 100


def test_error_convention_without_gpu():
    lib = _lib.load()
    rc = lib.r2d2_tree_layout(0, 8, None, None, None, None)
    assert rc == -1 and b"M must be positive" in lib.r2d2_last_error()
    with pytest.raises(RuntimeError, match="tree_layout"):
        _lib.call("r2d2_tree_layout", -5, 8, None, None, None, None)


@pytest.mark.parametrize("M", [1, 31, 32, 33, 1024, 1025, 500000, 4000000, 1 << 22])
def test_tree_layout_matches_oracle(M):
    from oracle.replay_oracle import SumTree
    sizes = (ctypes.c_int64 * 8)()
    offs = (ctypes.c_int64 * 8)()
    n = ctypes.c_int(0)
    tot = ctypes.c_int64(0)
    _lib.call("r2d2_tree_layout", M, 8, sizes, offs, ctypes.byref(n), ctypes.byref(tot))
    want = SumTree.level_sizes(M)
    assert list(sizes[: n.value]) == want
    assert tot.value == sum(want)
    assert list(offs[: n.value]) == list(np.cumsum([0] + want[:-1]))


@pytest.mark.parametrize("H", [256, 512])
def test_param_layout(H):
    lay = ParamLayout(H)
    total = sum(lay.numel(k) for k in MODULE_ORDER)
    assert lay.total == total
    assert total == {256: 2037095, 512: 4826727}[H]
    spans = sorted((lay.offset(k), lay.numel(k)) for k in MODULE_ORDER)
    pos = 0
    for off, n in spans:
        assert off == pos
        pos += n
    # tensors the kernels read as one block are adjacent
    assert lay.offset("adv.0.weight") == lay.offset("val.0.weight") + H * H
    assert lay.offset("adv.2.weight") == lay.offset("val.2.weight") + H
    assert lay.offset("adv.2.bias") == lay.offset("val.2.bias") + 1
    assert lay.offset("lstm.weight_hh") % 4 == 0 and lay.offset("lstm.weight_ih") % 4 == 0


def test_shard_bounds_cover():
    for total in (10, 2037095, 4826727):
        for world in (1, 2, 3, 8):
            b = [shard_bounds(total, world, r) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == total
            for (l0, h0), (l1, h1) in zip(b, b[1:]):
                assert h0 == l1
            assert all(lo % 4 == 0 or lo == total for lo, _ in b)


def test_no_product_import_of_oracle():
    pkg = os.path.join(ROOT, "pytorch_r2d2_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            assert "oracle" not in open(os.path.join(pkg, f)).read(), f


def test_dp_shard_elems_matches_library():
    """Host mirror of the optimiser-state shard size == the C entry point (no GPU work)."""
    from pytorch_r2d2_b200 import _lib, parallel
    lib = _lib.load()
    for n in (1, 4, 7, 100_003, 4_737_000):
        for world in (1, 2, 3, 8, 16):
            assert parallel.shard_elems(n, world) == lib.r2d2_dp_shard_elems(n, world)
            # the shards cover the float4 part and the tail fits behind the last one
            per = (n // 4 + world - 1) // world
            assert per * world >= n // 4 and parallel.shard_elems(n, world) >= per * 4 + n % 4
