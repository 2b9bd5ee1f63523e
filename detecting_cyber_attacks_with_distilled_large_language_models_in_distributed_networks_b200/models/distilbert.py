"""Pure-PyTorch ``DDoSClassifier`` with a state_dict identical to the reference's.

Reference: ``DDoSClassifier`` = HF ``DistilBertModel`` + ``Dropout(0.3)`` + ``Linear(768, 2)`` on the CLS token
(client1.py:53-65); the encoder body is transformers ``modeling_distilbert.py`` (:83-122 embeddings, :154-207
attention, :210-228 FFN, :231-263 block).  This module re-implements that body directly (no ``transformers``
dependency, no ``from_pretrained`` directory requirement — SURVEY.md Q14) with the same parameter names, so

* checkpoints move freely between the reference, this module and the B200 engine's arena;
* it is the numerical oracle for the CUDA engine and the model used by the CPU plumbing tier.

It is NOT the B200 hot path: on a GPU ``models/engine.py`` runs the hand-written kernels.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .arena import ModelDims


class Embeddings(nn.Module):
    def __init__(self, d: ModelDims):
        super().__init__()
        self.word_embeddings = nn.Embedding(d.vocab_size, d.dim)
        self.position_embeddings = nn.Embedding(d.max_position, d.dim)
        self.LayerNorm = nn.LayerNorm(d.dim, eps=d.ln_eps)
        self.dropout = nn.Dropout(d.dropout)

    def forward(self, input_ids: torch.Tensor) -> torch.Tensor:
        S = input_ids.shape[1]
        pos = torch.arange(S, device=input_ids.device)
        x = self.word_embeddings(input_ids) + self.position_embeddings(pos)[None]
        return self.dropout(self.LayerNorm(x))


class SelfAttention(nn.Module):
    def __init__(self, d: ModelDims):
        super().__init__()
        self.n_heads, self.dim = d.n_heads, d.dim
        self.q_lin = nn.Linear(d.dim, d.dim)
        self.k_lin = nn.Linear(d.dim, d.dim)
        self.v_lin = nn.Linear(d.dim, d.dim)
        self.out_lin = nn.Linear(d.dim, d.dim)
        self.dropout = nn.Dropout(d.attention_dropout)

    def forward(self, x: torch.Tensor, key_mask: Optional[torch.Tensor]) -> torch.Tensor:
        B, S, D = x.shape
        H, dh = self.n_heads, D // self.n_heads
        q = self.q_lin(x).view(B, S, H, dh).transpose(1, 2)
        k = self.k_lin(x).view(B, S, H, dh).transpose(1, 2)
        v = self.v_lin(x).view(B, S, H, dh).transpose(1, 2)
        scores = (q @ k.transpose(-1, -2)) / math.sqrt(dh)
        if key_mask is not None:  # [B, S] with 1 = attend
            scores = scores.masked_fill(~key_mask.bool()[:, None, None, :], torch.finfo(scores.dtype).min)
        w = self.dropout(torch.softmax(scores, dim=-1))
        ctx = (w @ v).transpose(1, 2).reshape(B, S, D)
        return self.out_lin(ctx)


class FFN(nn.Module):
    def __init__(self, d: ModelDims):
        super().__init__()
        self.lin1 = nn.Linear(d.dim, d.hidden_dim)
        self.lin2 = nn.Linear(d.hidden_dim, d.dim)
        self.dropout = nn.Dropout(d.dropout)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.dropout(self.lin2(F.gelu(self.lin1(x))))


class TransformerBlock(nn.Module):
    def __init__(self, d: ModelDims):
        super().__init__()
        self.attention = SelfAttention(d)
        self.sa_layer_norm = nn.LayerNorm(d.dim, eps=d.ln_eps)
        self.ffn = FFN(d)
        self.output_layer_norm = nn.LayerNorm(d.dim, eps=d.ln_eps)

    def forward(self, x: torch.Tensor, key_mask: Optional[torch.Tensor]) -> torch.Tensor:
        sa = self.sa_layer_norm(self.attention(x, key_mask) + x)
        return self.output_layer_norm(self.ffn(sa) + sa)


class Transformer(nn.Module):
    def __init__(self, d: ModelDims):
        super().__init__()
        self.layer = nn.ModuleList([TransformerBlock(d) for _ in range(d.n_layers)])

    def forward(self, x: torch.Tensor, key_mask: Optional[torch.Tensor]) -> torch.Tensor:
        for blk in self.layer:
            x = blk(x, key_mask)
        return x


class DistilBertEncoder(nn.Module):
    def __init__(self, d: ModelDims):
        super().__init__()
        self.embeddings = Embeddings(d)
        self.transformer = Transformer(d)

    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        return self.transformer(self.embeddings(input_ids), attention_mask)


class DDoSClassifier(nn.Module):
    """Binary DDoS classifier; ``forward(input_ids, attention_mask) -> logits [B, 2]`` (reference client1.py:60-65)."""

    def __init__(self, dims: ModelDims = ModelDims(), seed: Optional[int] = None):
        super().__init__()
        self.dims = dims
        self.distilbert = DistilBertEncoder(dims)
        self.dropout = nn.Dropout(dims.head_dropout)
        self.classifier = nn.Linear(dims.dim, dims.n_classes)
        if seed is not None:
            self.reset_parameters(seed)

    def reset_parameters(self, seed: int = 0) -> None:
        from .arena import ParamArena

        arena = ParamArena(self.dims).init_random_(seed)
        self.load_state_dict(arena.state_dict())

    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        h = self.distilbert(input_ids, attention_mask)
        return self.classifier(self.dropout(h[:, 0, :]))
