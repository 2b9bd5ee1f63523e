"""Load pretrained DistilBERT encoder weights from a Hugging Face model directory.

The reference fine-tunes ``DistilBertModel.from_pretrained('./distilbert-base-uncased')`` (client1.py:55-56, hard
failure at :357-361 when the directory is missing) under a freshly initialised ``Linear(768, 2)`` head.  This loader
gives every backend of this package the same starting point: it reads ``model.safetensors`` or ``pytorch_model.bin``
from the directory, maps the checkpoint's names onto the ``DDoSClassifier`` state_dict names (the encoder lives under
``distilbert.`` there; hub checkpoints of ``DistilBertForMaskedLM`` already carry that prefix and add ``vocab_*`` heads
that are dropped), and leaves ``classifier.*`` at its default initialisation.  When the directory holds no weights
(only a ``vocab.txt``, or nothing at all — this box has no network) the caller keeps its random initialisation and
says so in the log.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch

WEIGHT_FILES = ("model.safetensors", "pytorch_model.bin")
_DROP_PREFIXES = ("vocab_transform.", "vocab_layer_norm.", "vocab_projector.", "pre_classifier.", "classifier.",
                  "qa_outputs.", "dropout.")


def find_weights(model_dir: Optional[str]) -> Optional[str]:
    if not model_dir or not os.path.isdir(model_dir):
        return None
    for name in WEIGHT_FILES:
        p = os.path.join(model_dir, name)
        if os.path.isfile(p):
            return p
    return None


def read_checkpoint(path: str) -> Dict[str, torch.Tensor]:
    if path.endswith(".safetensors"):
        from safetensors.torch import load_file

        return load_file(path, device="cpu")
    return torch.load(path, map_location="cpu", weights_only=True)


def encoder_state_dict(ckpt: Dict[str, torch.Tensor], expected: Dict[str, Tuple[int, ...]]
                       ) -> "OrderedDict[str, torch.Tensor]":
    """Checkpoint tensors renamed to ``DDoSClassifier`` keys, restricted to encoder tensors whose shapes fit
    ``expected`` (name -> shape of the target model); everything else is ignored."""
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for k, v in ckpt.items():
        if k.startswith(_DROP_PREFIXES) or not isinstance(v, torch.Tensor) or not v.is_floating_point():
            continue
        name = k if k.startswith("distilbert.") else "distilbert." + k
        if name in expected and tuple(v.shape) == tuple(expected[name]):
            out[name] = v
    return out


def load_pretrained_encoder(backend, model_dir: Optional[str], log=None) -> int:
    """Loads the encoder weights found in ``model_dir`` into ``backend`` (anything with ``state_dict`` /
    ``load_state_dict``); returns the number of tensors loaded (0: none found, random initialisation kept)."""
    say = log.say if log is not None else (lambda m: None)
    path = find_weights(model_dir)
    if path is None:
        say(f"no pretrained weights in {model_dir!r} (looked for {', '.join(WEIGHT_FILES)}): training from random "
            f"initialisation — the reference fine-tunes pretrained DistilBERT, expect different metrics")
        return 0
    sd = backend.state_dict()
    expected = {k: tuple(v.shape) for k, v in sd.items()}
    enc = encoder_state_dict(read_checkpoint(path), expected)
    want = [k for k in expected if k.startswith("distilbert.")]
    if not enc:
        say(f"{path} holds no tensor that fits this model ({len(want)} encoder tensors expected): random "
            f"initialisation kept")
        return 0
    merged = OrderedDict((k, enc.get(k, v)) for k, v in sd.items())
    backend.load_state_dict(merged)
    missing = [k for k in want if k not in enc]
    say(f"loaded {len(enc)} of {len(want)} pretrained encoder tensors from {path}"
        + (f"; {len(missing)} kept their initialisation (first: {missing[0]})" if missing else "")
        + "; classifier head freshly initialised")
    return len(enc)
