"""NCCL all-reduce FedAvg — the BASELINE the fused peer-memory kernel is measured against (BASELINE.json: "a path
that only calls NCCL is the baseline, not the product").  ``ncclAllReduce(sum)`` over the flat arena followed by a
scale and a separate bf16 cast: three passes over memory where the product kernel does one."""
from __future__ import annotations

from typing import Optional

import torch


def nccl_fedavg_(arena: torch.Tensor, mirror: Optional[torch.Tensor] = None, mirror_elems: int = 0, group=None) -> None:
    import torch.distributed as dist

    dist.all_reduce(arena, op=dist.ReduceOp.SUM, group=group)
    arena.div_(dist.get_world_size(group))
    if mirror is not None and mirror_elems:
        mirror.copy_(arena[:mirror_elems])
