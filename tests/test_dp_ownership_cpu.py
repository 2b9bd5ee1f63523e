"""Ownership split of a data-parallel client (parallel/dp.py::_partition, models/engine.py::owned_optimizer_ranges):
on every group size the ranges the ranks own tile the parameter arena exactly once -- that is what makes the merged
training-state snapshot (sum of the owned ranges) equal the single-GPU layout.  Pure index logic: the objects are
stubs, no GPU memory is touched."""
import importlib
from types import SimpleNamespace

import numpy as np
import pytest

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
arena_mod = importlib.import_module(f"{PKG}.models.arena")


def _engine_cls():
    try:
        return importlib.import_module(f"{PKG}.models.engine").B200Engine
    except Exception as e:  # noqa: BLE001 - the native library is missing on this box
        pytest.skip(f"engine module not importable: {e}")


def _group(layout, rank, world):
    dp_mod = importlib.import_module(f"{PKG}.parallel.dp")
    g = dp_mod.DPGroup.__new__(dp_mod.DPGroup)          # index logic only: no peer buffers
    g.rank, g.world, g.layout = rank, world, layout
    g.gemm_bounds, g.r2_bounds = g._partition()
    return g


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("sparse_emb", [True, False])
@pytest.mark.parametrize("small", [True, False])
def test_owned_ranges_tile_the_arena_once(world, sparse_emb, small):
    eng_cls = _engine_cls()
    dims = (arena_mod.ModelDims(vocab_size=1000, max_position=128, dim=256, n_layers=2, n_heads=4, hidden_dim=512)
            if small else arena_mod.ModelDims())
    lay = arena_mod.ArenaLayout(dims)
    cover = np.zeros(lay.total, dtype=np.int8)
    per_rank = []
    for r in range(world):
        g = _group(lay, r, world)
        stub = SimpleNamespace(arena=SimpleNamespace(layout=lay), dp=g, dp_sparse_emb=sparse_emb, dims=dims)
        ranges = eng_cls.owned_optimizer_ranges(stub)
        assert all(0 <= lo < hi <= lay.total and lo % 4 == 0 and hi % 4 == 0 for lo, hi in ranges)
        for lo, hi in ranges:
            cover[lo:hi] += 1
        per_rank.append(sum(hi - lo for lo, hi in ranges))
        # the GEMM weights a rank owns are whole tensors (a weight-gradient GEMM reduce-adds into ONE arena)
        for lo, hi in g.gemm_ranges(r):
            assert g.gemm_owner(lo) == r and g.gemm_owner(hi - 1) == r
    assert int(cover.min()) == 1 and int(cover.max()) == 1, "every element has exactly one owner"
    assert sum(per_rank) == lay.total
    if not small:   # the point of the interleaved split: no rank carries much more optimizer work than the others
        assert max(per_rank) < 1.6 * lay.total / world


def test_single_gpu_owns_everything():
    eng_cls = _engine_cls()
    lay = arena_mod.ArenaLayout(arena_mod.ModelDims())
    stub = SimpleNamespace(arena=SimpleNamespace(layout=lay), dp=None)
    assert eng_cls.owned_optimizer_ranges(stub) == [(0, lay.total)]
