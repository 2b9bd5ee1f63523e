"""Intra-client data parallelism (BASELINE.json config 4): one federated client spans G GPUs.

The reference has no data parallelism at all; this is the B200-native addition.  Every GPU of the
group runs the same step engine on its own micro-batch.  Communication is fused into the compute kernels instead of
being separate NCCL collectives:

* gradient reduce-scatter — each weight-gradient GEMM stores its output tile with a TMA reduce-add straight into
  the gradient arena of the GPU that OWNS that weight (``cp.reduce.async.bulk.tensor`` on a peer-mapped address), tile by
  tile, overlapping the rest of backward.  The small atomically-accumulated gradients (biases, LayerNorm,
  embeddings, classifier) are reduce-scattered by one peer-read kernel after backward.
* sharded Adam — each GPU updates only the arena elements it owns (ZeRO-1/2-like: optimizer state and reduced
  gradients are not replicated).
* parameter all-gather — fused into the optimizer kernel: the updated bf16 GEMM weights are stored into every peer's
  mirror, the updated fp32 values that kernels read directly (embeddings, LayerNorm, biases) into every peer's arena.
* two one-block cross-GPU barriers per step (signal pads), replayable from a CUDA graph.

``torch.distributed`` is used only to exchange CUDA IPC handles at set-up.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from ..models.arena import ArenaLayout
from ..ops import native as nv

PAD_BYTES = 64 * 4


class DPGroup:
    """Buffers and ownership map of one data-parallel client group.

    ``members`` are the global ranks of this client's GPUs (this process must be one of them).  IPC handles are
    exchanged over ``ipc_group`` (default: the world), so every rank maps every other rank's buffers once; the DP
    kernels use the members' subset, the inter-client FedAvg uses all of them.
    """

    def __init__(self, rank: int, world: int, layout: ArenaLayout, dist_group=None, members: Optional[List[int]] = None,
                 member_group=None):
        import torch.distributed as dist

        self.global_rank, self.global_world = rank, world
        self.members = list(range(world)) if members is None else list(members)
        if rank not in self.members:
            raise ValueError("this rank is not a member of its own client group")
        if len(self.members) > 8:
            raise ValueError("a data-parallel client group spans at most 8 GPUs")
        self.rank, self.world, self.layout = self.members.index(rank), len(self.members), layout
        self._dist_group = member_group if member_group is not None else dist_group
        n = layout.total
        self.param_buf = nv.RawDeviceBuffer(nv.p2p_alloc(n * 4), n * 4)
        self.grad_buf = nv.RawDeviceBuffer(nv.p2p_alloc(n * 4), n * 4)
        self.mirror_buf = nv.RawDeviceBuffer(nv.p2p_alloc(layout.gemm_end * 2), layout.gemm_end * 2)
        self.pad_buf = nv.RawDeviceBuffer(nv.p2p_alloc(PAD_BYTES), PAD_BYTES)
        # which vocabulary rows this GPU has touched (embed_ln_bwd writes it); peers read it in the row-sparse optimizer
        vocab_bytes = (layout.dims.vocab_size + 255) // 256 * 256
        self.flags_buf = nv.RawDeviceBuffer(nv.p2p_alloc(vocab_bytes), vocab_bytes)
        mine = {k: nv.p2p_ipc_get(b.ptr) for k, b in (("param", self.param_buf), ("grad", self.grad_buf),
                                                      ("mirror", self.mirror_buf), ("pad", self.pad_buf),
                                                      ("flags", self.flags_buf))}
        gathered: List[Optional[dict]] = [None] * world
        dist.all_gather_object(gathered, mine, group=dist_group)
        self._opened: List[nv.RawDeviceBuffer] = []
        self.all_param_ptrs, self.all_grad_ptrs, self.all_mirror_ptrs, self.all_pad_ptrs = [], [], [], []
        self.all_flag_ptrs: List[int] = []
        for r, h in enumerate(gathered):
            for name, lst, own in (("param", self.all_param_ptrs, self.param_buf),
                                   ("grad", self.all_grad_ptrs, self.grad_buf),
                                   ("mirror", self.all_mirror_ptrs, self.mirror_buf),
                                   ("pad", self.all_pad_ptrs, self.pad_buf),
                                   ("flags", self.all_flag_ptrs, self.flags_buf)):
                if r == rank:
                    lst.append(own.ptr)
                elif name in ("grad", "flags") and r not in self.members:
                    lst.append(0)  # gradients (and row flags) are only ever exchanged inside a client group
                else:
                    p = nv.p2p_ipc_open(h[name])
                    self._opened.append(nv.RawDeviceBuffer(p, 0, owner=False, ipc=True))
                    lst.append(p)
        pick = lambda lst: [lst[g] for g in self.members]  # noqa: E731
        self.param_ptrs, self.grad_ptrs = pick(self.all_param_ptrs), pick(self.all_grad_ptrs)
        self.mirror_ptrs, self.pad_ptrs = pick(self.all_mirror_ptrs), pick(self.all_pad_ptrs)
        self.flag_ptrs = pick(self.all_flag_ptrs)
        self._peer_grad_bufs = [nv.RawDeviceBuffer(p, n * 4, owner=False) for p in self.grad_ptrs]
        self.gemm_bounds, self.r2_bounds = self._partition()
        self._fed_epoch = 0
        self._world_group = dist_group
        dist.barrier(group=dist_group)

    @classmethod
    def build(cls, rank: int, world: int, gpus_per_client: int, layout: ArenaLayout) -> "DPGroup":
        """Collective over the world: creates one process group per client and returns this rank's DPGroup."""
        import torch.distributed as dist

        if world % gpus_per_client:
            raise ValueError("world size must be a multiple of gpus_per_client")
        mine_members, mine_group = None, None
        for c in range(world // gpus_per_client):
            members = list(range(c * gpus_per_client, (c + 1) * gpus_per_client))
            g = dist.new_group(members)  # every rank must take part in every new_group call
            if rank in members:
                mine_members, mine_group = members, g
        return cls(rank, world, layout, dist_group=None, members=mine_members, member_group=mine_group)

    @property
    def num_clients(self) -> int:
        return self.global_world // self.world

    def fedavg_clients(self, weights=None, max_ctas: int = 0, stream=None) -> None:
        """Inter-client FedAvg when clients span several GPUs: the first GPU of every client contributes its (complete)
        fp32 arena, all GPUs of all clients receive the average and its bf16 copy.  Call
        ``allgather_fp32_gemm_params()`` first so the contributed arenas are complete."""
        self._fed_epoch += 1
        reps = [self.all_param_ptrs[c * self.world] for c in range(self.num_clients)]
        nv.fedavg_p2p_grouped(reps, self.all_param_ptrs, self.all_mirror_ptrs, self.all_pad_ptrs, self.global_rank,
                              self.layout.total, self._fed_epoch, weights=weights, bf16_begin=0,
                              bf16_end=self.layout.gemm_end, max_ctas=max_ctas, stream=stream)

    # ---- ownership -----------------------------------------------------------------------------------------
    def _gemm_units(self) -> List[Tuple[int, int]]:
        """(offset, numel) of every GEMM weight in arena order, q|k|v fused into one unit (one wgrad GEMM each)."""
        lay, d = self.layout, self.layout.dims
        units = []
        for l in range(d.n_layers):
            p = f"distilbert.transformer.layer.{l}."
            units.append((lay.qkv_weight_offset(l), 3 * d.dim * d.dim))
            for name in ("attention.out_lin.weight", "ffn.lin1.weight", "ffn.lin2.weight"):
                units.append((lay.offsets[p + name], lay.numel(p + name)))
        return units

    def _partition(self):
        """Ownership of the GEMM weights and of the fp32 remainder of the arena.

        A weight-gradient GEMM reduce-adds its whole output into ONE arena, so GEMM tensors are owned whole.  The four
        tensors of a transformer block go to four different ranks (largest first: lin1, lin2, fused qkv, out_lin), the
        same slots in every block — for groups of 8 the blocks alternate between two sets of four ranks.  Each block's
        optimizer work is then spread over the group instead of falling to the one or two ranks that would own the
        block in a contiguous split: with per-block early Adam (models/engine.py) the update of the LAST block of the
        backward pass is the exposed one, and it shrinks from a whole block to its largest tensor (a third of it)."""
        lay, G = self.layout, self.world
        units = self._gemm_units()                       # per layer, arena order: qkv, out_lin, lin1, lin2
        slot_of = {0: 2, 1: 3, 2: 0, 3: 1}
        owned: List[List[Tuple[int, int]]] = [[] for _ in range(G)]
        self._unit_owner: List[Tuple[int, int, int]] = []
        for i, (off, n) in enumerate(units):
            layer, j = divmod(i, 4)
            r = (slot_of[j] + 4 * layer) % G
            owned[r].append((off, off + n))
            self._unit_owner.append((off, off + n, r))
        self.gemm_owned = owned
        r2 = lay.total - lay.gemm_end
        r2b = [lay.gemm_end + (r2 * r // G) // 4 * 4 for r in range(G)] + [lay.total]
        return None, r2b

    def gemm_ranges(self, r: int) -> List[Tuple[int, int]]:
        """Arena element ranges [lo, hi) of the GEMM weights rank ``r`` owns (whole tensors, arena order)."""
        return self.gemm_owned[r]

    def r2_range(self, r: int) -> Tuple[int, int]:
        return self.r2_bounds[r], self.r2_bounds[r + 1]

    def gemm_owner(self, offset: int) -> int:
        for lo, hi, r in self._unit_owner:
            if lo <= offset < hi:
                return r
        raise ValueError(offset)

    # ---- views ---------------------------------------------------------------------------------------------
    def param_tensor(self) -> torch.Tensor:
        return self.param_buf.view(torch.float32, numel=self.layout.total)

    def grad_tensor(self) -> torch.Tensor:
        return self.grad_buf.view(torch.float32, numel=self.layout.total)

    def mirror_tensor(self) -> torch.Tensor:
        return self.mirror_buf.view(torch.bfloat16, numel=self.layout.gemm_end)

    def pad_tensor(self) -> torch.Tensor:
        return self.pad_buf.view(torch.int32, numel=64)

    def flags_tensor(self) -> torch.Tensor:
        """uint8 [vocab]: rows of the word-embedding table this GPU has touched (peer-readable)."""
        return self.flags_buf.view(torch.uint8, numel=self.layout.dims.vocab_size)

    def word_rows(self, r: int) -> Tuple[int, int]:
        """Vocabulary rows whose optimizer state rank ``r`` owns (row-sparse embedding update)."""
        V = self.layout.dims.vocab_size
        return V * r // self.world, V * (r + 1) // self.world

    def self_first(self, lst: List[int]) -> List[int]:
        """``lst`` (one entry per member) rotated so that this rank's entry comes first."""
        return lst[self.rank:] + lst[: self.rank]

    def owner_grad_view(self, offset: int, shape: Tuple[int, ...]) -> torch.Tensor:
        """View of the gradient of the GEMM weight at ``offset`` inside its OWNER's gradient arena (peer memory when
        the owner is another GPU): the destination of the wgrad TMA reduce-add."""
        owner = self.gemm_owner(offset)
        numel = 1
        for s in shape:
            numel *= s
        return self._peer_grad_bufs[owner].view(torch.float32, offset_bytes=offset * 4, numel=numel).view(shape)

    # ---- collectives used outside the step -------------------------------------------------------------------
    def allgather_fp32_gemm_params(self, stream=None) -> None:
        """After training: every rank pushes the fp32 master values of the GEMM weights it owns to all peers, so that
        each rank holds a complete fp32 ``state_dict`` (checkpoints, inter-client FedAvg)."""
        import torch.distributed as dist

        for lo, hi in self.gemm_ranges(self.rank):
            for k in range(self.world):
                if k != self.rank:
                    nv.fedavg_local([self.param_ptrs[self.rank] + lo * 4], self.param_ptrs[k] + lo * 4, hi - lo,
                                    weights=[1.0], stream=stream)
        torch.cuda.current_stream().synchronize()
        dist.barrier(group=self._dist_group)

    def reset_barrier_flags(self) -> None:
        """Host-synchronised clear of the barrier words (needed whenever the device step counter is rewound)."""
        import torch.distributed as dist

        torch.cuda.synchronize()
        dist.barrier(group=self._dist_group)
        self.pad_tensor()[48:64].zero_()
        torch.cuda.synchronize()
        dist.barrier(group=self._dist_group)

    def error_flag(self, clear: bool = True) -> int:
        """1 when a bounded handshake wait of a kernel on this rank expired since the flag was last cleared.  Reading
        it reports the failure ONCE: the word is reset so that later rounds are judged on their own."""
        pad = self.pad_tensor()
        v = int(pad[33].item())
        if v and clear:
            pad[33] = 0
        return v

    def close(self) -> None:
        for b in self._opened:
            b.free()
        self._opened = []
        for b in (self.param_buf, self.grad_buf, self.mirror_buf, self.pad_buf, self.flags_buf):
            b.free()
