"""The C++ batch tokenizer (csrc/host/tokenizer.cpp) must agree with the Python WordPiece implementation — which is
itself checked against HF's DistilBertTokenizer in test_data_cpu.py — on every input, including the rows it declines
(non-ASCII text falls back to Python)."""
from __future__ import annotations

import os
import random
import string

import numpy as np
import pytest

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.data import (
    cicids, synth, tokenizer as tk)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.utils import build


@pytest.fixture(scope="module")
def native_built():
    build.build_host()
    assert tk._load_native() is not None
    return True


def _python_encode(tok, texts, max_len):
    ids = np.full((len(texts), max_len), tok.pad_id, dtype=np.int32)
    lens = np.zeros(len(texts), dtype=np.int32)
    for i, t in enumerate(texts):
        row, n = tok.encode(t, max_len)
        ids[i], lens[i] = row, n
    return ids, lens


def _corpus():
    rnd = random.Random(0)
    df = cicids.clean_frame(synth.synth_flows(600, seed=3))
    texts = df.apply(cicids.features_to_text, axis=1).tolist()
    alphabet = string.ascii_letters + string.digits + string.punctuation + "   \t\n\r\x00\x01\x0b\x1f\x7f"
    for _ in range(600):
        texts.append("".join(rnd.choice(alphabet) for _ in range(rnd.randint(0, 300))))
    texts += ["", " ", "a" * 101 + " ok", "x" * 100, "Destination PORT is 443!!", "1234567890" * 40,
              "café résumé naïve", "中文 mixed 123", "tab\tsep\nnew\rline", "[CLS] [SEP] [UNK]"]
    return texts


@pytest.mark.parametrize("max_len", [128, 16, 3])
def test_native_matches_python(native_built, max_len):
    tok = tk.WordPieceTokenizer(tk.synthetic_vocab())
    assert tok._native_handle() is not None
    texts = _corpus()
    ids_n, lens_n = tok.encode_batch(texts, max_len)
    ids_p, lens_p = _python_encode(tok, texts, max_len)
    bad = np.nonzero((ids_n != ids_p).any(axis=1) | (lens_n != lens_p))[0]
    assert bad.size == 0, (bad[:5], [texts[i][:60] for i in bad[:3]])


def test_native_with_a_real_style_vocab_and_switch_off(native_built, tmp_path, monkeypatch):
    vocab = ["[PAD]"] + [f"[unused{i}]" for i in range(99)] + ["[UNK]", "[CLS]", "[SEP]", "[MASK]"]
    vocab += list("abcdefghijklmnopqrstuvwxyz0123456789.:,") + ["##" + c for c in "abcdefghijklmnopqrstuvwxyz0123456789"]
    vocab += ["port", "flow", "##s", "dest", "##ination", "is", "12", "##34"]
    path = tk.write_vocab(str(tmp_path / "vocab.txt"), vocab)
    tok = tk.WordPieceTokenizer.from_vocab_file(path)
    texts = ["Destination ports is 1234.", "flows: 12,34 zzz ü", "UNKNOWN_§ token"]
    a = tok.encode_batch(texts, 32)
    monkeypatch.setenv("DCAD_NATIVE_TOK", "0")
    tok2 = tk.WordPieceTokenizer.from_vocab_file(path)
    assert tok2._native_handle() is None
    b = tok2.encode_batch(texts, 32)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert a[0][0, 0] == 101 and a[0][0, a[1][0] - 1] == 102
