"""Full training-state checkpoint (utils/checkpoint.py): resuming reproduces the uninterrupted run bit for bit on the
Torch backend (model + Adam moments + RNG + shuffle order), which the reference's model-only .pth cannot."""
from __future__ import annotations

import torch

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.data.pack import (
    PackedSplit)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.models.arena import (
    ModelDims)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.train.trainer import (
    TorchBackend)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.utils import (
    checkpoint as ck)


def _dims():
    return ModelDims(vocab_size=200, max_position=32, dim=64, n_layers=1, n_heads=1, hidden_dim=128)


def _split(n=24, S=16, vocab=200, seed=0):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(1, vocab, (n, S), generator=g, dtype=torch.int32)
    lens = torch.randint(4, S + 1, (n,), generator=g, dtype=torch.int32)
    labels = torch.randint(0, 2, (n,), generator=g, dtype=torch.int32)
    return PackedSplit(ids, lens, labels)


def test_resume_is_bit_exact_on_torch_backend(tmp_path):
    split = _split()
    a = TorchBackend(_dims(), batch=8, seq=16, lr=1e-3, seed=3, init_seed=1)
    gen = torch.Generator().manual_seed(7)
    a.train_epoch(split, gen)
    path = ck.save_training_state(str(tmp_path / "state.pt"), a, round_idx=2, epoch=1, shuffle_generator=gen,
                                  extra={"note": "after epoch 1"})
    a.train_epoch(split, gen)                       # uninterrupted continuation
    want = {k: v.clone() for k, v in a.state_dict().items()}

    b = TorchBackend(_dims(), batch=8, seq=16, lr=1e-3, seed=99, init_seed=5)   # different init, different RNG
    gen_b = torch.Generator().manual_seed(12345)
    prog = ck.load_training_state(path, b, shuffle_generator=gen_b)
    assert prog["round"] == 2 and prog["epoch"] == 1 and prog["note"] == "after epoch 1"
    b.train_epoch(split, gen_b)
    got = b.state_dict()
    assert all(torch.equal(got[k], want[k]) for k in want)

    # the model-only checkpoint of the reference loses the Adam moments: same weights, different continuation
    c = TorchBackend(_dims(), batch=8, seq=16, lr=1e-3, seed=3, init_seed=1)
    st = torch.load(path, weights_only=False)
    c.load_state_dict(st["model"])
    gen_c = torch.Generator()
    gen_c.set_state(st["progress"]["shuffle_generator"])
    torch.set_rng_state(st["rng"]["torch_cpu"])
    c.train_epoch(split, gen_c)
    assert any(not torch.equal(c.state_dict()[k], want[k]) for k in want)


def test_rejects_foreign_snapshots(tmp_path):
    a = TorchBackend(_dims(), batch=8, seq=16)
    st = ck.training_state(a)
    st["backend"] = "b200"
    try:
        ck.load_training_state(st, a)
        raise AssertionError("expected a ValueError")
    except ValueError:
        pass


def test_non_writing_participant_leaves_no_file(tmp_path):
    """``write=False`` is what every rank but the reporter of a data-parallel client passes: it takes part in the
    snapshot (on GPUs: the collectives that merge the sharded Adam moments) and writes nothing."""
    a = TorchBackend(_dims(), batch=8, seq=16)
    path = str(tmp_path / "state.pt")
    assert ck.save_training_state(path, a, write=False, round_idx=1) == path
    assert list(tmp_path.iterdir()) == []
    ck.save_training_state(path, a, round_idx=1)
    assert [p.name for p in tmp_path.iterdir()] == ["state.pt"]   # atomic write: no temp file left behind
