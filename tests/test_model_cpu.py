"""Model contract: state_dict keys/shapes identical to the reference's DDoSClassifier (HF DistilBertModel body) and
forward parity with HF on CPU."""
from __future__ import annotations

import pytest
import torch

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.models.arena import (
    ArenaLayout, ModelDims, ParamArena, state_dict_keys)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.models.distilbert import (
    DDoSClassifier)


def test_state_dict_contract_full_size():
    keys = state_dict_keys(ModelDims())
    assert len(keys) == 102
    assert sum(torch.Size(s).numel() for _, s in keys) == 66_364_418
    lay = ArenaLayout(ModelDims())
    assert lay.num_params == 66_364_418 and lay.total % 4096 == 0
    # q|k|v weights are contiguous so the fused [2304,768] view is free
    o = lay.offsets
    p = "distilbert.transformer.layer.3.attention."
    assert o[p + "k_lin.weight"] == o[p + "q_lin.weight"] + 768 * 768
    assert o[p + "v_lin.weight"] == o[p + "k_lin.weight"] + 768 * 768
    assert o[p + "k_lin.bias"] == o[p + "q_lin.bias"] + 768


def _small():
    return ModelDims(vocab_size=500, max_position=64, dim=64, n_layers=2, n_heads=4, hidden_dim=128)


def test_module_keys_match_arena_and_hf():
    transformers = pytest.importorskip("transformers")
    d = _small()
    m = DDoSClassifier(d, seed=1)
    assert [k for k, _ in m.state_dict().items()] == [k for k, _ in state_dict_keys(d)]
    cfg = transformers.DistilBertConfig(vocab_size=d.vocab_size, max_position_embeddings=d.max_position, dim=d.dim,
                                        n_layers=d.n_layers, n_heads=d.n_heads, hidden_dim=d.hidden_dim)
    hf = transformers.DistilBertModel(cfg)
    hf_keys = ["distilbert." + k for k in hf.state_dict().keys()] + ["classifier.weight", "classifier.bias"]
    assert hf_keys == [k for k, _ in state_dict_keys(d)]


def test_forward_matches_hf_distilbert():
    transformers = pytest.importorskip("transformers")
    d = _small()
    m = DDoSClassifier(d, seed=2).eval()
    cfg = transformers.DistilBertConfig(vocab_size=d.vocab_size, max_position_embeddings=d.max_position, dim=d.dim,
                                        n_layers=d.n_layers, n_heads=d.n_heads, hidden_dim=d.hidden_dim)
    hf = transformers.DistilBertModel(cfg).eval()
    hf.load_state_dict({k[len("distilbert."):]: v for k, v in m.state_dict().items() if k.startswith("distilbert.")})
    ids = torch.randint(0, d.vocab_size, (3, 32))
    mask = torch.ones(3, 32, dtype=torch.long)
    mask[1, 20:] = 0
    mask[2, 5:] = 0
    with torch.no_grad():
        ours = m.distilbert(ids, mask)
        theirs = hf(input_ids=ids, attention_mask=mask)[0]
        logits = m(ids, mask)
    valid = mask.bool()
    assert torch.allclose(ours[valid], theirs[valid], atol=1e-5, rtol=1e-4)
    assert torch.allclose(logits, theirs[:, 0, :] @ m.classifier.weight.t() + m.classifier.bias, atol=1e-5)


def test_arena_roundtrip_and_reference_checkpoint_format(tmp_path):
    d = _small()
    a = ParamArena(d).init_random_(3)
    sd = a.state_dict(clone=True)
    path = tmp_path / "client1_model.pth"
    torch.save(sd, path)                       # same format the reference writes (client1.py:388)
    m = DDoSClassifier(d)
    m.load_state_dict(torch.load(path))        # loads into a module...
    b = ParamArena(d)
    b.load_state_dict(m.state_dict())          # ...and back into an arena
    assert torch.equal(a.flat, b.flat)
    with pytest.raises(KeyError):
        b.load_state_dict({"bogus": torch.zeros(1)})


def test_pretrained_encoder_loader(tmp_path):
    """An HF DistilBertModel saved with save_pretrained() (what ./distilbert-base-uncased is for the reference,
    client1.py:55-56) lands in the encoder of the DDoSClassifier; the head keeps its own initialisation; a directory
    without weights leaves the model alone."""
    from transformers import DistilBertConfig, DistilBertModel

    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.models import (
        arena as A, distilbert as D, pretrained as P)

    dims = A.ModelDims(vocab_size=120, max_position=32, dim=64, n_heads=2, hidden_dim=128, n_layers=2)
    hf = DistilBertModel(DistilBertConfig(vocab_size=120, max_position_embeddings=32, dim=64, n_heads=2,
                                          hidden_dim=128, n_layers=2))
    for ext_dir, safe in (("st", True), ("bin", False)):
        d = tmp_path / ext_dir
        hf.save_pretrained(d, safe_serialization=safe)
        model = D.DDoSClassifier(dims)
        head_before = model.state_dict()["classifier.weight"].clone()
        said = []
        n = P.load_pretrained_encoder(model, str(d), type("L", (), {"say": staticmethod(said.append)})())
        sd = model.state_dict()
        assert n == len(hf.state_dict()) == len(sd) - 2, said
        for k, v in hf.state_dict().items():
            assert torch.equal(sd["distilbert." + k], v), k
        assert torch.equal(sd["classifier.weight"], head_before)
    empty = tmp_path / "empty"
    empty.mkdir()
    model = D.DDoSClassifier(dims)
    before = {k: v.clone() for k, v in model.state_dict().items()}
    assert P.load_pretrained_encoder(model, str(empty)) == 0
    assert all(torch.equal(v, model.state_dict()[k]) for k, v in before.items())
