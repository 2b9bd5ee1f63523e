"""Peer-memory runtime for the NVSwitch FedAvg path (one process per GPU, or one process driving several GPUs).

The reference moves each client's 265 MB state_dict through pickle + gzip + TCP twice per round (client1.py:276-336,
server.py:57-114; ≈31–43 s per round).  On an NVSwitch box every client's parameter arena instead lives
in peer-mappable device memory, and aggregation is one kernel per GPU that loads the peers' arenas and stores the
average (and the bf16 compute copy) into every client's buffers (``csrc/fedavg.cu``).  NCCL is not involved:
``torch.distributed`` is used only to exchange the 64-byte CUDA IPC handles at set-up.

``PeerGroup``       multi-process: cudaMalloc'd arena / bf16 mirror / signal pad per rank, IPC handles all-gathered.
``LocalPeerGroup``  single process, N devices: direct peer access (used by tests and the single-process launcher).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..ops import native as nv

PAD_BYTES = 64 * 4
TWO_SHOT, ONE_SHOT, NVLS = 0, 1, 2


class PeerGroup:
    def __init__(self, rank: int, world: int, arena_elems: int, mirror_elems: int, dist_group=None,
                 with_out_buffer: bool = False, arena_dtype: torch.dtype = torch.float32):
        import torch.distributed as dist

        self.rank, self.world = rank, world
        self.arena_elems, self.mirror_elems = arena_elems, mirror_elems
        self.arena_dtype = arena_dtype
        esz = 4 if arena_dtype == torch.float32 else 2
        self.arena_buf = nv.RawDeviceBuffer(nv.p2p_alloc(arena_elems * esz), arena_elems * esz)
        self.mirror_buf = nv.RawDeviceBuffer(nv.p2p_alloc(max(mirror_elems, 8) * 2), max(mirror_elems, 8) * 2)
        self.pad_buf = nv.RawDeviceBuffer(nv.p2p_alloc(PAD_BYTES), PAD_BYTES)
        self.out_buf = (nv.RawDeviceBuffer(nv.p2p_alloc(arena_elems * esz), arena_elems * esz)
                        if with_out_buffer else None)
        mine = {"arena": nv.p2p_ipc_get(self.arena_buf.ptr), "mirror": nv.p2p_ipc_get(self.mirror_buf.ptr),
                "pad": nv.p2p_ipc_get(self.pad_buf.ptr),
                "out": nv.p2p_ipc_get(self.out_buf.ptr) if self.out_buf else None}
        gathered: List[Optional[dict]] = [None] * world
        dist.all_gather_object(gathered, mine, group=dist_group)
        self._opened: List[nv.RawDeviceBuffer] = []
        self.arena_ptrs, self.mirror_ptrs, self.pad_ptrs, self.out_ptrs = [], [], [], []
        for r, h in enumerate(gathered):
            if r == rank:
                self.arena_ptrs.append(self.arena_buf.ptr)
                self.mirror_ptrs.append(self.mirror_buf.ptr)
                self.pad_ptrs.append(self.pad_buf.ptr)
                self.out_ptrs.append(self.out_buf.ptr if self.out_buf else 0)
            else:
                for name, lst in (("arena", self.arena_ptrs), ("mirror", self.mirror_ptrs), ("pad", self.pad_ptrs),
                                  ("out", self.out_ptrs)):
                    if h[name] is None:
                        lst.append(0)
                        continue
                    p = nv.p2p_ipc_open(h[name])
                    self._opened.append(nv.RawDeviceBuffer(p, 0, owner=False, ipc=True))
                    lst.append(p)
        self.epoch = 0
        dist.barrier(group=dist_group)

    # views for the engine
    def arena_tensor(self) -> torch.Tensor:
        return self.arena_buf.view(self.arena_dtype, numel=self.arena_elems)

    def mirror_tensor(self) -> torch.Tensor:
        return self.mirror_buf.view(torch.bfloat16, numel=self.mirror_elems)

    def out_tensor(self) -> Optional[torch.Tensor]:
        return self.out_buf.view(self.arena_dtype, numel=self.arena_elems) if self.out_buf else None

    def pad_tensor(self) -> torch.Tensor:
        return self.pad_buf.view(torch.int32, numel=64)

    def fedavg(self, mode: int = TWO_SHOT, weights: Optional[Sequence[float]] = None,
               participants: Optional[Sequence[int]] = None, with_mirror: bool = True, max_ctas: int = 0,
               stream=None, epoch: Optional[int] = None) -> None:
        """Launch this rank's share of the aggregation (all participating ranks must call it).  In place for two-shot;
        one-shot writes the ``out`` buffer.  ``participants`` restricts the average to a subset (quorum).  ``epoch``
        (strictly increasing, the same on every participant — e.g. the round number + 1) replaces the group's own
        launch counter, which drifts apart when a client sits a round out."""
        ranks = list(range(self.world)) if participants is None else list(participants)
        if self.rank not in ranks:
            return
        if mode == NVLS:
            mode = TWO_SHOT  # (this group has no multicast mapping)
        self.epoch = self.epoch + 1 if epoch is None else int(epoch)
        sel = lambda lst: [lst[r] for r in ranks]  # noqa: E731
        nv.fedavg_p2p(sel(self.arena_ptrs), sel(self.pad_ptrs), ranks.index(self.rank), self.arena_elems, self.epoch,
                      out_ptrs=sel(self.out_ptrs) if mode == ONE_SHOT else None,
                      mirror_ptrs=sel(self.mirror_ptrs) if (with_mirror and self.mirror_elems) else None,
                      weights=weights, in_bf16=self.arena_dtype == torch.bfloat16, mode=mode, bf16_begin=0,
                      bf16_end=self.mirror_elems if with_mirror else 0, max_ctas=max_ctas, stream=stream)

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
        for b in (self.arena_buf, self.mirror_buf, self.pad_buf, self.out_buf):
            if b is not None:
                b.free()


class SymmPeerGroup:
    """Peer group whose buffers come from ``torch.distributed._symmetric_memory`` (used purely as an allocator and
    handle exchange): one symmetric allocation per rank laid out as ``[fp32 arena | bf16 mirror | signal pad]``.  When
    the host exposes NVLink multicast the handle carries a multicast address and ``fedavg(NVLS)`` reduces inside the
    NVSwitch; the unicast peer pointers serve the two-shot kernel otherwise (or on request)."""

    def __init__(self, rank: int, world: int, arena_elems: int, mirror_elems: int, dist_group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm

        self.rank, self.world = rank, world
        self.arena_elems, self.mirror_elems = arena_elems, mirror_elems
        self.arena_dtype = torch.float32
        self._arena_bytes = arena_elems * 4
        self._mirror_off = (self._arena_bytes + 255) // 256 * 256
        self._pad_off = (self._mirror_off + max(mirror_elems, 8) * 2 + 255) // 256 * 256
        total = self._pad_off + PAD_BYTES
        dev = torch.device("cuda", torch.cuda.current_device())
        self._buf = symm.empty(total, dtype=torch.uint8, device=dev)
        self._buf.zero_()
        group = dist_group if dist_group is not None else dist.group.WORLD
        self._hdl = symm.rendezvous(self._buf, group=group)
        base = [int(p) for p in self._hdl.buffer_ptrs]
        self.arena_ptrs = base
        self.mirror_ptrs = [p + self._mirror_off for p in base]
        self.pad_ptrs = [p + self._pad_off for p in base]
        mc = int(self._hdl.multicast_ptr) if getattr(self._hdl, "has_multicast_support", True) else 0
        self.multicast_ptr = mc
        self.epoch = 0
        torch.cuda.synchronize()
        dist.barrier(group=dist_group)

    @property
    def has_nvls(self) -> bool:
        return self.multicast_ptr != 0

    def arena_tensor(self) -> torch.Tensor:
        return self._buf[: self._arena_bytes].view(torch.float32)

    def mirror_tensor(self) -> torch.Tensor:
        return self._buf[self._mirror_off: self._mirror_off + self.mirror_elems * 2].view(torch.bfloat16)

    def pad_tensor(self) -> torch.Tensor:
        return self._buf[self._pad_off: self._pad_off + PAD_BYTES].view(torch.int32)

    def out_tensor(self):
        return None

    def fedavg(self, mode: int = NVLS, weights: Optional[Sequence[float]] = None, with_mirror: bool = True,
               max_ctas: int = 0, stream=None, participants: Optional[Sequence[int]] = None,
               epoch: Optional[int] = None, local_cast: bool = False) -> None:
        """``mode`` NVLS: in-switch reduction (all ranks, unweighted); it falls back to the two-shot kernel over the
        unicast peer pointers when the host has no multicast, when per-client weights are given, or when only a
        subset of the clients takes part (the multicast mapping always spans every rank).

        ``local_cast`` (NVLS only): the kernel averages the fp32 arena alone and every GPU then derives its bf16
        compute copy from its own arena with a local cast kernel, instead of every rank multicasting the bf16 copy of
        its shard to all others -- (N-1)/N x 85 MB less NVLink ingress per GPU for 255 MB of local HBM traffic."""
        ranks = list(range(self.world)) if participants is None else list(participants)
        if self.rank not in ranks:
            return
        self.epoch = self.epoch + 1 if epoch is None else int(epoch)
        full = len(ranks) == self.world
        if mode == NVLS and self.has_nvls and weights is None and full:
            bcast_mirror = with_mirror and not local_cast
            nv.fedavg_nvls(self.multicast_ptr, self.multicast_ptr + self._mirror_off if bcast_mirror else 0,
                           self.pad_ptrs, self.rank, self.arena_elems, self.epoch, bf16_begin=0,
                           bf16_end=self.mirror_elems if bcast_mirror else 0, max_ctas=max_ctas, stream=stream)
            if with_mirror and local_cast and self.mirror_elems:
                # stream-ordered behind the kernel, whose end handshake means every rank's shard has landed here
                nv.cast_f32_bf16(self.arena_tensor()[: self.mirror_elems], self.mirror_tensor(), stream=stream)
        else:
            sel = lambda lst: [lst[r] for r in ranks]  # noqa: E731
            nv.fedavg_p2p(sel(self.arena_ptrs), sel(self.pad_ptrs), ranks.index(self.rank), self.arena_elems, self.epoch,
                          mirror_ptrs=sel(self.mirror_ptrs) if (with_mirror and self.mirror_elems) else None,
                          weights=weights, mode=TWO_SHOT, bf16_begin=0,
                          bf16_end=self.mirror_elems if with_mirror else 0, max_ctas=max_ctas, stream=stream)

    def error_flag(self, clear: bool = True) -> int:
        """1 when a bounded handshake wait of a kernel on this rank expired since the flag was last cleared.  Reading
        it reports the failure ONCE: the word is reset so that later rounds are judged on their own."""
        pad = self.pad_tensor()
        v = int(pad[33].item())
        if v and clear:
            pad[33] = 0
        return v

    def close(self) -> None:
        self._hdl = None
        self._buf = None


class LocalPeerGroup:
    """All ranks in one process (one CUDA device each) with direct peer access."""

    def __init__(self, devices: Sequence[int], arena_elems: int, mirror_elems: int, with_out_buffer: bool = False,
                 arena_dtype: torch.dtype = torch.float32):
        self.devices = list(devices)
        self.world = len(self.devices)
        self.arena_elems, self.mirror_elems, self.arena_dtype = arena_elems, mirror_elems, arena_dtype
        esz = 4 if arena_dtype == torch.float32 else 2
        self.arena_bufs, self.mirror_bufs, self.pad_bufs, self.out_bufs = [], [], [], []
        for d in self.devices:
            with torch.cuda.device(d):
                for peer in self.devices:
                    if peer != d:
                        if not nv.p2p_can_access(d, peer):
                            raise nv.NativeError(f"device {d} cannot access peer {peer}")
                        nv.p2p_enable_peer_access(peer)
                self.arena_bufs.append(nv.RawDeviceBuffer(nv.p2p_alloc(arena_elems * esz), arena_elems * esz))
                self.mirror_bufs.append(nv.RawDeviceBuffer(nv.p2p_alloc(max(mirror_elems, 8) * 2), max(mirror_elems, 8) * 2))
                self.pad_bufs.append(nv.RawDeviceBuffer(nv.p2p_alloc(PAD_BYTES), PAD_BYTES))
                if with_out_buffer:
                    self.out_bufs.append(nv.RawDeviceBuffer(nv.p2p_alloc(arena_elems * esz), arena_elems * esz))
        self.epoch = 0

    def arena_tensor(self, r: int) -> torch.Tensor:
        with torch.cuda.device(self.devices[r]):
            return self.arena_bufs[r].view(self.arena_dtype, numel=self.arena_elems)

    def mirror_tensor(self, r: int) -> torch.Tensor:
        with torch.cuda.device(self.devices[r]):
            return self.mirror_bufs[r].view(torch.bfloat16, numel=self.mirror_elems)

    def out_tensor(self, r: int) -> torch.Tensor:
        with torch.cuda.device(self.devices[r]):
            return self.out_bufs[r].view(self.arena_dtype, numel=self.arena_elems)

    def fedavg_all(self, mode: int = TWO_SHOT, weights: Optional[Sequence[float]] = None, with_mirror: bool = True,
                   max_ctas: int = 0) -> None:
        """Launch every rank's kernel (each on its own device's current stream); the kernels synchronise among
        themselves through the signal pads."""
        self.epoch += 1
        for r, d in enumerate(self.devices):
            with torch.cuda.device(d):
                nv.fedavg_p2p([b.ptr for b in self.arena_bufs], [b.ptr for b in self.pad_bufs], r, self.arena_elems,
                              self.epoch, out_ptrs=[b.ptr for b in self.out_bufs] if mode == ONE_SHOT else None,
                              mirror_ptrs=[b.ptr for b in self.mirror_bufs] if (with_mirror and self.mirror_elems) else None,
                              weights=weights, in_bf16=self.arena_dtype == torch.bfloat16, mode=mode, bf16_begin=0,
                              bf16_end=self.mirror_elems if with_mirror else 0, max_ctas=max_ctas)

    def close(self) -> None:
        for lst in (self.arena_bufs, self.mirror_bufs, self.pad_bufs, self.out_bufs):
            for r, b in enumerate(lst):
                with torch.cuda.device(self.devices[r]):
                    b.free()
