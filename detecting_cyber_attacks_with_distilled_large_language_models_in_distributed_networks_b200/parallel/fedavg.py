"""FedAvg over state_dicts and over flat arenas.

Reference arithmetic (server.py:72-76): ``base = models[0]``; for every key ``base[k] += models[i][k]`` for
i = 1..N-1 in order, then ``base[k] /= N`` — an UNWEIGHTED mean over ALL keys.  Both paths below
reproduce it bit for bit (fp32 accumulate in client order, then a true division):

* ``fedavg_state_dicts``  — server-side entry point used by the socket fallback and by tests.  With a GPU the
  state_dicts are flattened into arenas and averaged by the hand-written kernel (one launch); without one the
  reference's loop runs on the CPU.  Non-float tensors (none in today's HF layout) are copied from client 0 instead
  of crashing like the reference does.
* ``fedavg_arenas_local`` — N arenas reachable from one device -> destination arena (+ bf16 mirror), kernel only.

Extensions: optional per-client weights (``n_k / n`` weighting) and a participant subset (quorum) — the caller
simply passes the subset.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, List, Optional, Sequence

import torch

from ..ops import native as nv


def fedavg_state_dicts_cpu(models: Sequence[Dict[str, torch.Tensor]],
                           weights: Optional[Sequence[float]] = None) -> "OrderedDict[str, torch.Tensor]":
    """The reference's loop, out of place (the reference mutates client 0's dict)."""
    n = len(models)
    if n == 0:
        raise ValueError("no models to aggregate")
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for k, v0 in models[0].items():
        if not isinstance(v0, torch.Tensor) or not v0.is_floating_point():
            out[k] = v0.clone() if isinstance(v0, torch.Tensor) else v0
            continue
        if weights is None:
            acc = v0.detach().to("cpu", copy=True)
            for m in models[1:]:
                acc += m[k].detach().to("cpu")
            acc /= n
        else:
            acc = v0.detach().to("cpu", copy=True) * float(weights[0])
            for w, m in zip(weights[1:], models[1:]):
                acc += float(w) * m[k].detach().to("cpu")
        out[k] = acc
    if hasattr(models[0], "_metadata"):
        out._metadata = models[0]._metadata  # type: ignore[attr-defined]
    return out


def _flatten(sd: Dict[str, torch.Tensor], keys: List[str], device) -> torch.Tensor:
    total = sum(sd[k].numel() for k in keys)
    pad = (-total) % 4
    flat = torch.empty(total + pad, dtype=torch.float32, device=device)
    off = 0
    for k in keys:
        n = sd[k].numel()
        flat[off:off + n].copy_(sd[k].reshape(-1), non_blocking=True)
        off += n
    if pad:
        flat[off:].zero_()
    return flat


def fedavg_state_dicts(models: Sequence[Dict[str, torch.Tensor]], weights: Optional[Sequence[float]] = None,
                       device: Optional[torch.device] = None,
                       out_device: str = "cpu") -> "OrderedDict[str, torch.Tensor]":
    """Average N state_dicts.  GPU kernel when available, otherwise the CPU loop."""
    if not nv.available():
        return fedavg_state_dicts_cpu(models, weights)
    device = device or torch.device("cuda", torch.cuda.current_device())
    fkeys = [k for k, v in models[0].items() if isinstance(v, torch.Tensor) and v.dtype == torch.float32]
    flats = [_flatten(m, fkeys, device) for m in models]
    dst = torch.empty_like(flats[0])
    nv.fedavg_local([f.data_ptr() for f in flats], dst.data_ptr(), dst.numel(), weights=weights)
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    off = 0
    host = dst.to(out_device) if out_device != str(device) else dst
    for k, v0 in models[0].items():
        if k in fkeys and isinstance(v0, torch.Tensor) and v0.dtype == torch.float32:
            n = v0.numel()
            out[k] = host[off:off + n].view(v0.shape).clone()
            off += n
        elif isinstance(v0, torch.Tensor) and v0.is_floating_point():  # other float dtypes: CPU arithmetic
            out[k] = fedavg_state_dicts_cpu([{k: m[k]} for m in models], weights)[k]
        else:
            out[k] = v0.clone() if isinstance(v0, torch.Tensor) else v0
    if hasattr(models[0], "_metadata"):
        out._metadata = models[0]._metadata  # type: ignore[attr-defined]
    return out


def fedavg_arenas_local(srcs: Sequence[torch.Tensor], dst: torch.Tensor, dst_bf16: Optional[torch.Tensor] = None,
                        bf16_end: int = 0, weights: Optional[Sequence[float]] = None, stream=None) -> torch.Tensor:
    """dst = mean(srcs) for flat arenas of equal length on (or peer-mapped into) the current device."""
    n = dst.numel()
    in_bf16 = srcs[0].dtype == torch.bfloat16
    assert all(s.numel() == n and s.dtype == srcs[0].dtype for s in srcs) and dst.dtype == srcs[0].dtype
    nv.fedavg_local([s.data_ptr() for s in srcs], dst.data_ptr(), n, in_bf16=in_bf16, weights=weights,
                    dst_bf16_ptr=dst_bf16.data_ptr() if dst_bf16 is not None else 0, bf16_begin=0,
                    bf16_end=bf16_end if dst_bf16 is not None else 0, stream=stream)
    return dst
