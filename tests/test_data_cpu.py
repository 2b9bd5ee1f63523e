"""L0 data layer parity with the reference (client1.py:68-93,365-366) and tokenizer parity with HF."""
from __future__ import annotations

import numpy as np
import pandas as pd
import pytest

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.data import (
    cicids, pack, split, synth, tokenizer)


@pytest.fixture(scope="module")
def csv_path(tmp_path_factory):
    p = tmp_path_factory.mktemp("data") / "CICIDS2017.csv"
    synth.write_synthetic_csv(str(p), 800, seed=3)
    return str(p)


def test_synthetic_csv_has_reference_schema(csv_path):
    header = open(csv_path).readline().rstrip("\n").split(",")
    assert len(header) == 79 and header.count("Fwd Header Length") == 2 and header[-1] == "Label"
    assert " Flow IAT Max" in header  # leading-blank headers preserved
    raw = open(csv_path).read()
    assert "Infinity" in raw
    df = pd.read_csv(csv_path)
    assert set(df["Label"].unique()) == {"BENIGN", "DDoS"}
    assert np.isinf(df["Flow Bytes/s"]).sum() >= 1


def test_features_to_text_matches_reference_bytes(csv_path, ref_client1):
    df = pd.read_csv(csv_path)
    df.replace([np.inf, -np.inf], np.nan, inplace=True)
    df.fillna(df.mean(numeric_only=True), inplace=True)
    expect = df.apply(ref_client1.features_to_text, axis=1).tolist()
    assert cicids.frame_to_texts(df) == expect
    assert cicids.features_to_text(df.iloc[0]) == expect[0]


@pytest.mark.parametrize("client_id", [1, 2])
def test_preprocess_data_matches_reference(csv_path, client_id, monkeypatch):
    from conftest import load_reference_module

    ref = load_reference_module(f"client{client_id}")
    rt, rl = ref.preprocess_data(csv_path, 0.25)
    texts, labels = cicids.preprocess_data(csv_path, 0.25, cicids.client_seed(client_id))
    assert texts == rt and labels == rl
    assert cicids.client_seed(1) == 42 and cicids.client_seed(2) == 43


def test_split_is_the_reference_60_20_20():
    from sklearn.model_selection import train_test_split

    texts = [f"t{i}" for i in range(503)]
    labels = list(range(503))
    xtr, xtmp, ytr, ytmp = train_test_split(texts, labels, test_size=0.4, random_state=42)
    xva, xte, yva, yte = train_test_split(xtmp, ytmp, test_size=0.5, random_state=42)
    got = split.split_lists(texts, labels, 42)
    assert got == (xtr, xva, xte, ytr, yva, yte)
    tr, va, te = split.split_indices(503, 42)
    assert len(tr) == 301 and len(va) == 101 and len(te) == 101


def test_disjoint_shards_do_not_overlap(csv_path):
    a, _ = cicids.preprocess_data(csv_path, 1.0, 42, shard=(0, 2))
    b, _ = cicids.preprocess_data(csv_path, 1.0, 43, shard=(1, 2))
    assert len(a) + len(b) == 800


def test_wordpiece_matches_hf_tokenizer(tmp_path):
    transformers = pytest.importorskip("transformers")
    vocab = tokenizer.synthetic_vocab()
    assert len(vocab) == 30522 and vocab[0] == "[PAD]" and vocab[100] == "[UNK]" and vocab[101] == "[CLS]"
    assert vocab[102] == "[SEP]" and vocab[103] == "[MASK]" and len(set(vocab)) == 30522
    d = tmp_path / "distilbert-base-uncased"
    tokenizer.write_vocab(str(d / "vocab.txt"), vocab)
    hf = transformers.DistilBertTokenizer.from_pretrained(str(d))
    ours = tokenizer.WordPieceTokenizer.from_dir_or_synthetic(str(d))
    texts = cicids.frame_to_texts(cicids.clean_frame(synth.synth_flows(40, seed=5)))
    texts += ["Héllo, wörld!  unknownwordzzz 12345678901234567890 a/b", ""]
    for t in texts:
        enc = hf(t, add_special_tokens=True, max_length=128, padding="max_length", truncation=True)
        ids, n = ours.encode(t, 128)
        assert ids == enc["input_ids"], t
        assert n == sum(enc["attention_mask"])


def test_pack_and_iterate():
    tok = tokenizer.WordPieceTokenizer(tokenizer.synthetic_vocab())
    texts = cicids.frame_to_texts(cicids.clean_frame(synth.synth_flows(50, seed=1)))
    ps = pack.pack_texts(tok, texts, [i % 2 for i in range(50)])
    assert ps.input_ids.shape == (50, 128) and int(ps.lens.max()) <= 128
    assert (ps.attention_mask().sum(1) == ps.lens).all()
    assert ps.num_batches(16) == 4
    sizes = [b[0].shape[0] for b in pack.iterate_batches(ps, 16)]
    assert sizes == [16, 16, 16, 2]
    import torch
    g1 = torch.Generator().manual_seed(7)
    g2 = torch.Generator().manual_seed(7)
    a = [b[2].tolist() for b in pack.iterate_batches(ps, 16, shuffle=True, generator=g1)]
    b = [b[2].tolist() for b in pack.iterate_batches(ps, 16, shuffle=True, generator=g2)]
    assert a == b  # seeded shuffling is reproducible (the reference's is not: SURVEY Q12)


def test_staged_epoch_matches_iterate_batches_and_dp_slices():
    """The once-per-epoch gather feeds the same batches in the same order as the per-step gather; a data-parallel
    client's GPUs share one permutation, take disjoint slices and keep the ragged end as an equal smaller step."""
    import torch

    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.data.pack import (
        PackedSplit, iterate_batches, stage_shuffled_epoch)

    n, S, B = 70, 8, 16
    split = PackedSplit(torch.arange(n * S, dtype=torch.int32).view(n, S), torch.arange(n, dtype=torch.int32) % S + 1,
                        torch.arange(n, dtype=torch.int32) % 2)
    want = list(iterate_batches(split, B, shuffle=True, generator=torch.Generator().manual_seed(5)))
    st = stage_shuffled_epoch(split, B, torch.Generator().manual_seed(5))
    assert len(st) == len(want) == 5 and st.steps.shape == (4, B * S + 2 * B)
    for i in range(4):
        rec = st.steps[i]
        assert torch.equal(rec[: B * S].view(B, S), want[i][0])
        assert torch.equal(rec[B * S: B * S + B], want[i][1]) and torch.equal(rec[B * S + B:], want[i][2])
    assert all(torch.equal(a, b) for a, b in zip(st.tail, want[4]))
    # two GPUs of one client: 70 = 2 full global steps of 32 + 6 left -> a last step of 3 samples each
    parts = [stage_shuffled_epoch(split, B, torch.Generator().manual_seed(5), rank=r, world=2) for r in range(2)]
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(5))
    seen = []
    for r, p in enumerate(parts):
        assert p.steps.shape[0] == 2 and p.tail[0].shape[0] == 3
        for i in range(2):
            ids = p.steps[i][: B * S].view(B, S)
            assert torch.equal(ids, split.input_ids[perm[i * 32 + r * B: i * 32 + (r + 1) * B]])
            seen += (ids[:, 0] // S).tolist()
        seen += (p.tail[0][:, 0] // S).tolist()
    assert len(set(seen)) == len(seen) == 70


def test_trimming_all_padding_columns_is_exact():
    """Columns that are padding in every sample can be dropped: same logits from the PyTorch module (the oracle of the
    engine) on [B, 128] inputs and on the same inputs trimmed to the executed length."""
    import torch

    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.data.pack import (
        PackedSplit, executed_seq_len)
    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.models import (
        arena as A, distilbert as D)

    assert executed_seq_len(128, 88) == 96 and executed_seq_len(128, 96) == 96 and executed_seq_len(128, 97) == 128
    assert executed_seq_len(128, 500) == 128 and executed_seq_len(64, 3) == 32
    g = torch.Generator().manual_seed(0)
    n, S = 6, 128
    lens = torch.tensor([88, 80, 79, 85, 12, 83], dtype=torch.int32)
    ids = torch.randint(5, 900, (n, S), generator=g, dtype=torch.int32)
    ids[torch.arange(S)[None, :] >= lens[:, None]] = 0
    split = PackedSplit(ids, lens, torch.zeros(n, dtype=torch.int32))
    seq = executed_seq_len(S, split.max_len())
    short = split.trim(seq)
    assert short.seq_len == 96 and torch.equal(short.input_ids, ids[:, :96])
    try:
        split.trim(64)
        raise AssertionError("trimming below the longest sample must be refused")
    except ValueError:
        pass
    dims = A.ModelDims(vocab_size=1000, max_position=128, dim=64, n_heads=2, hidden_dim=128, n_layers=2)
    model = D.DDoSClassifier(dims).eval()
    with torch.no_grad():
        full = model(split.input_ids.long(), split.attention_mask())
        cut = model(short.input_ids.long(), short.attention_mask())
    assert torch.allclose(full, cut, atol=1e-5, rtol=1e-5)
