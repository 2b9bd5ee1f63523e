"""Flat parameter arena with an HF-compatible ``state_dict`` view.

The reference model's interchange format is the plain ``OrderedDict`` returned by ``DDoSClassifier.state_dict()``
(reference client1.py:53-65,388; SURVEY.md Appendix C: 102 fp32 tensors, 66,364,418 parameters, keys
``distilbert.*`` and ``classifier.{weight,bias}``).  Here all 102 tensors are views into ONE contiguous fp32
buffer so that

* Adam is one launch over the arena (``adam_flat``),
* FedAvg is one kernel over peer arenas (no per-key Python loop, reference server.py:72-76),
* checkpoints stay byte-compatible with the reference (``state_dict()`` yields the same keys/shapes/order).

Arena order (NOT state_dict order): first every GEMM weight (q|k|v adjacent so the fused ``[2304,768]`` QKV weight
is a plain view), then biases / LayerNorm / classifier / position and word embeddings.  The first region is the
one mirrored in bf16 for the tensor cores; the second region holds every gradient that kernels accumulate with
atomics and therefore must be cleared each step.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, Iterator, List, Optional, Tuple

import torch

ALIGN_ELEMS = 64  # every tensor starts on a 256-byte boundary (TMA needs 16 B, float4 paths need 16 B)
TOTAL_ALIGN = 4096


@dataclass(frozen=True)
class ModelDims:
    vocab_size: int = 30522
    max_position: int = 512
    dim: int = 768
    n_layers: int = 6
    n_heads: int = 12
    hidden_dim: int = 3072
    n_classes: int = 2
    ln_eps: float = 1e-12
    dropout: float = 0.1
    attention_dropout: float = 0.1
    head_dropout: float = 0.3
    initializer_range: float = 0.02

    @property
    def head_dim(self) -> int:
        return self.dim // self.n_heads


def state_dict_keys(d: ModelDims) -> List[Tuple[str, Tuple[int, ...]]]:
    """Keys and shapes in the exact order of the reference's ``state_dict``."""
    out: List[Tuple[str, Tuple[int, ...]]] = [
        ("distilbert.embeddings.word_embeddings.weight", (d.vocab_size, d.dim)),
        ("distilbert.embeddings.position_embeddings.weight", (d.max_position, d.dim)),
        ("distilbert.embeddings.LayerNorm.weight", (d.dim,)),
        ("distilbert.embeddings.LayerNorm.bias", (d.dim,)),
    ]
    for l in range(d.n_layers):
        p = f"distilbert.transformer.layer.{l}."
        for lin in ("q_lin", "k_lin", "v_lin", "out_lin"):
            out.append((p + f"attention.{lin}.weight", (d.dim, d.dim)))
            out.append((p + f"attention.{lin}.bias", (d.dim,)))
        out.append((p + "sa_layer_norm.weight", (d.dim,)))
        out.append((p + "sa_layer_norm.bias", (d.dim,)))
        out.append((p + "ffn.lin1.weight", (d.hidden_dim, d.dim)))
        out.append((p + "ffn.lin1.bias", (d.hidden_dim,)))
        out.append((p + "ffn.lin2.weight", (d.dim, d.hidden_dim)))
        out.append((p + "ffn.lin2.bias", (d.dim,)))
        out.append((p + "output_layer_norm.weight", (d.dim,)))
        out.append((p + "output_layer_norm.bias", (d.dim,)))
    out.append(("classifier.weight", (d.n_classes, d.dim)))
    out.append(("classifier.bias", (d.n_classes,)))
    return out


def _is_gemm_weight(key: str) -> bool:
    return key.endswith(".weight") and ("_lin.weight" in key or "ffn.lin" in key)


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class ArenaLayout:
    """Offsets of every state_dict tensor inside the flat arena."""

    def __init__(self, dims: ModelDims):
        self.dims = dims
        self.keys = state_dict_keys(dims)
        self.shapes: Dict[str, Tuple[int, ...]] = dict(self.keys)
        self.offsets: Dict[str, int] = {}
        off = 0
        # region 1: GEMM weights, layer by layer, q|k|v|out|lin1|lin2 — q,k,v contiguous (768*768 is 64-aligned)
        for l in range(dims.n_layers):
            p = f"distilbert.transformer.layer.{l}."
            for name in ("attention.q_lin.weight", "attention.k_lin.weight", "attention.v_lin.weight",
                         "attention.out_lin.weight", "ffn.lin1.weight", "ffn.lin2.weight"):
                k = p + name
                self.offsets[k] = off
                n = 1
                for s in self.shapes[k]:
                    n *= s
                assert n % ALIGN_ELEMS == 0, "GEMM weights must keep q|k|v contiguous"
                off += n
        self.gemm_end = off
        # region 2: per-layer biases with q|k|v contiguous too (fused [2304] bias), then the rest
        for l in range(dims.n_layers):
            p = f"distilbert.transformer.layer.{l}."
            for name in ("attention.q_lin.bias", "attention.k_lin.bias", "attention.v_lin.bias"):
                self.offsets[p + name] = off
                off += self.shapes[p + name][0]
            off = _round_up(off, ALIGN_ELEMS)
        for k, shp in self.keys:
            if k in self.offsets:
                continue
            n = 1
            for s in shp:
                n *= s
            self.offsets[k] = off
            off = _round_up(off + n, ALIGN_ELEMS)
        self.used = off
        self.total = _round_up(off, TOTAL_ALIGN)
        self.num_params = sum(int(torch.Size(s).numel()) for _, s in self.keys)

    def numel(self, key: str) -> int:
        return int(torch.Size(self.shapes[key]).numel())

    def qkv_weight_offset(self, layer: int) -> int:
        return self.offsets[f"distilbert.transformer.layer.{layer}.attention.q_lin.weight"]

    def qkv_bias_offset(self, layer: int) -> int:
        return self.offsets[f"distilbert.transformer.layer.{layer}.attention.q_lin.bias"]


class ParamArena:
    """One flat fp32 buffer (+ optional same-layout grad / Adam buffers and a bf16 mirror of the GEMM weights)."""

    def __init__(self, dims: ModelDims = ModelDims(), device: str | torch.device = "cpu",
                 flat: Optional[torch.Tensor] = None, with_optimizer_state: bool = False, with_bf16: bool = False):
        self.dims = dims
        self.layout = ArenaLayout(dims)
        self.device = torch.device(device)
        n = self.layout.total
        if flat is None:
            flat = torch.zeros(n, dtype=torch.float32, device=self.device)
        assert flat.numel() == n and flat.dtype == torch.float32 and flat.is_contiguous()
        self.flat = flat
        self.grad: Optional[torch.Tensor] = None
        self.exp_avg: Optional[torch.Tensor] = None
        self.exp_avg_sq: Optional[torch.Tensor] = None
        self.bf16: Optional[torch.Tensor] = None
        if with_optimizer_state:
            self.grad = torch.zeros_like(flat)
            self.exp_avg = torch.zeros_like(flat)
            self.exp_avg_sq = torch.zeros_like(flat)
        if with_bf16:
            self.bf16 = torch.zeros(self.layout.gemm_end, dtype=torch.bfloat16, device=flat.device)

    # ---- views -------------------------------------------------------------------------------------------
    def view(self, key: str, buf: Optional[torch.Tensor] = None) -> torch.Tensor:
        buf = self.flat if buf is None else buf
        off = self.layout.offsets[key]
        return buf[off: off + self.layout.numel(key)].view(self.layout.shapes[key])

    def bf16_view(self, key: str) -> torch.Tensor:
        assert self.bf16 is not None and self.layout.offsets[key] < self.layout.gemm_end
        off = self.layout.offsets[key]
        return self.bf16[off: off + self.layout.numel(key)].view(self.layout.shapes[key])

    def qkv_weight(self, layer: int, buf: Optional[torch.Tensor] = None) -> torch.Tensor:
        buf = self.flat if buf is None else buf
        d = self.dims.dim
        off = self.layout.qkv_weight_offset(layer)
        return buf[off: off + 3 * d * d].view(3 * d, d)

    def qkv_bias(self, layer: int, buf: Optional[torch.Tensor] = None) -> torch.Tensor:
        buf = self.flat if buf is None else buf
        off = self.layout.qkv_bias_offset(layer)
        return buf[off: off + 3 * self.dims.dim]

    def items(self) -> Iterator[Tuple[str, torch.Tensor]]:
        for k, _ in self.layout.keys:
            yield k, self.view(k)

    # ---- state_dict interchange ---------------------------------------------------------------------------
    def state_dict(self, clone: bool = False) -> "OrderedDict[str, torch.Tensor]":
        sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        for k, t in self.items():
            sd[k] = t.clone() if clone else t
        return sd

    def load_state_dict(self, sd: Dict[str, torch.Tensor], strict: bool = True) -> None:
        missing = [k for k, _ in self.layout.keys if k not in sd]
        unexpected = [k for k in sd if k not in self.layout.shapes]
        if strict and (missing or unexpected):
            raise KeyError(f"state_dict mismatch: missing={missing[:4]}... unexpected={unexpected[:4]}...")
        with torch.no_grad():
            for k, _ in self.layout.keys:
                if k in sd:
                    src = sd[k]
                    if tuple(src.shape) != self.layout.shapes[k]:
                        raise ValueError(f"shape mismatch for {k}: {tuple(src.shape)} vs {self.layout.shapes[k]}")
                    if src.is_floating_point():
                        self.view(k).copy_(src.to(dtype=torch.float32))
                    # integer / bool buffers (none in the current HF layout) are ignored: SURVEY.md Q3
        self.refresh_bf16()

    def refresh_bf16(self) -> None:
        if self.bf16 is not None:
            self.bf16.copy_(self.flat[: self.layout.gemm_end])

    # ---- init ---------------------------------------------------------------------------------------------
    def init_random_(self, seed: int = 0) -> "ParamArena":
        """HF DistilBERT ``_init_weights``: N(0, 0.02) for Linear/Embedding weights, zeros for biases, LN = (1, 0);
        the classifier uses nn.Linear's default init like the reference (client1.py:58)."""
        g = torch.Generator(device="cpu").manual_seed(seed)
        d = self.dims
        with torch.no_grad():
            self.flat.zero_()
            for k, shp in self.layout.keys:
                v = self.view(k)
                if k.startswith("classifier."):
                    bound = 1.0 / (d.dim ** 0.5)
                    v.copy_((torch.rand(shp, generator=g) * 2 - 1) * bound)
                elif "LayerNorm.weight" in k or "layer_norm.weight" in k:
                    v.fill_(1.0)
                elif k.endswith(".bias"):
                    v.zero_()
                else:
                    v.copy_(torch.randn(shp, generator=g) * d.initializer_range)
        self.refresh_bf16()
        return self
