"""FedAvg arithmetic parity with the reference server (server.py:67-79)."""
from __future__ import annotations

import copy
from collections import OrderedDict

import pytest
import torch

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.models.arena import (
    ModelDims, ParamArena)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel.fedavg import (
    fedavg_state_dicts, fedavg_state_dicts_cpu)


def _models(n, seed=0):
    d = ModelDims(vocab_size=300, max_position=32, dim=64, n_layers=1, n_heads=4, hidden_dim=128)
    return [ParamArena(d).init_random_(seed + i).state_dict(clone=True) for i in range(n)]


@pytest.mark.parametrize("n", [2, 3, 5])
def test_matches_reference_aggregate_models(ref_server, n, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)  # the reference writes ddos_distilbert_model.pth into the cwd
    models = _models(n)
    ref_server.NUM_CLIENTS = n
    ref_server.received_models.clear()
    ref_server.received_models.extend(copy.deepcopy(models))
    expect = ref_server.aggregate_models()
    got = fedavg_state_dicts_cpu(models)
    assert list(got.keys()) == list(expect.keys())
    for k in expect:
        assert torch.equal(got[k], expect[k]), k       # bit parity: same order of fp32 adds, true division
    got2 = fedavg_state_dicts(models)                   # dispatcher (CPU path here, GPU kernel on a GPU box)
    for k in expect:
        assert torch.allclose(got2[k], expect[k], rtol=0, atol=0), k
    ref_server.received_models.clear()


def test_inputs_are_not_mutated_and_non_float_buffers_survive():
    models = _models(2)
    for m in models:
        m["position_ids"] = torch.arange(8)           # the reference crashes on this
    snap = copy.deepcopy(models[0])
    out = fedavg_state_dicts_cpu(models)
    assert torch.equal(out["position_ids"], torch.arange(8))
    for k in snap:
        assert torch.equal(models[0][k], snap[k])


def test_weighted_average():
    models = _models(3)
    out = fedavg_state_dicts_cpu(models, weights=[0.5, 0.25, 0.25])
    k = "classifier.weight"
    exp = 0.5 * models[0][k] + 0.25 * models[1][k] + 0.25 * models[2][k]
    assert torch.allclose(out[k], exp, atol=1e-7)
