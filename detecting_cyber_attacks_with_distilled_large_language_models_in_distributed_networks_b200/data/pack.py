"""Device-resident, pre-tokenised datasets and batch iteration.

Replaces the reference's ``CICIDS2017Dataset`` + ``DataLoader(batch_size=16, shuffle=True, num_workers=0)``
(client1.py:26-50,367-372): every split is tokenised once, stored as ``input_ids [N, S] int32``, ``lens [N] int32``
and ``labels [N] int32`` on the training device, and epochs iterate over an on-device permutation drawn from a seeded
generator (the reference's shuffle is unseeded).  The last batch may be smaller, as with
``drop_last=False``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterator, Optional, Sequence, Tuple

import numpy as np
import torch


@dataclass
class PackedSplit:
    input_ids: torch.Tensor  # [N, S] int32
    lens: torch.Tensor       # [N] int32
    labels: torch.Tensor     # [N] int32

    def __len__(self) -> int:
        return int(self.input_ids.shape[0])

    @property
    def seq_len(self) -> int:
        return int(self.input_ids.shape[1])

    def to(self, device) -> "PackedSplit":
        return PackedSplit(self.input_ids.to(device), self.lens.to(device), self.labels.to(device))

    def pin(self) -> "PackedSplit":
        return PackedSplit(self.input_ids.pin_memory(), self.lens.pin_memory(), self.labels.pin_memory())

    def subset(self, idx) -> "PackedSplit":
        idx = torch.as_tensor(idx, dtype=torch.long, device=self.input_ids.device)
        return PackedSplit(self.input_ids[idx], self.lens[idx], self.labels[idx])

    def attention_mask(self) -> torch.Tensor:
        S = self.seq_len
        return (torch.arange(S, device=self.lens.device)[None, :] < self.lens[:, None]).to(torch.int64)

    def num_batches(self, batch_size: int) -> int:
        return (len(self) + batch_size - 1) // batch_size

    def max_len(self) -> int:
        return int(self.lens.max().item()) if len(self) else 0

    def trim(self, seq: int) -> "PackedSplit":
        """Drops the trailing columns past ``seq`` (every sample must be at most ``seq`` tokens long)."""
        if seq >= self.seq_len:
            return self
        if self.max_len() > seq:
            raise ValueError(f"cannot trim to {seq}: a sample has {self.max_len()} tokens")
        return PackedSplit(self.input_ids[:, :seq].contiguous(), self.lens, self.labels)


def executed_seq_len(max_len: int, longest: int, multiple: int = 32) -> int:
    """Sequence length the step actually runs at: the longest tokenised sample rounded up to ``multiple``, at most
    ``max_len``.  The reference pads every sample to ``max_length=128`` (client1.py:38-45) and computes every padded
    position; a padded position influences nothing (masked as a key, never read as an output, zero gradient), so
    columns that are padding in EVERY sample can be dropped without changing a single result -- flow sentences are
    78-88 tokens, i.e. 96 columns instead of 128: a quarter fewer token rows in every GEMM / LayerNorm / attention."""
    longest = max(int(longest), 1)
    return min(int(max_len), (longest + multiple - 1) // multiple * multiple)


def pack_texts(tokenizer, texts: Sequence[str], labels: Sequence[int], max_len: int = 128) -> PackedSplit:
    ids, lens = tokenizer.encode_batch(texts, max_len)
    return PackedSplit(torch.from_numpy(ids), torch.from_numpy(lens),
                       torch.as_tensor(np.asarray(labels, dtype=np.int32)))


def iterate_batches(split: PackedSplit, batch_size: int, shuffle: bool = False,
                    generator: Optional[torch.Generator] = None,
                    drop_last: bool = False) -> Iterator[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]]:
    n = len(split)
    dev = split.input_ids.device
    if shuffle:
        if generator is not None and generator.device != dev:
            perm = torch.randperm(n, generator=generator).to(dev)
        else:
            perm = torch.randperm(n, generator=generator, device=dev)
    else:
        perm = None
    for s in range(0, n, batch_size):
        e = min(n, s + batch_size)
        if drop_last and e - s < batch_size:
            break
        if perm is None:
            yield split.input_ids[s:e], split.lens[s:e], split.labels[s:e]
        else:
            ix = perm[s:e]
            yield split.input_ids[ix], split.lens[ix], split.labels[ix]


@dataclass
class StagedEpoch:
    """One shuffled epoch laid out for the step engine: ``steps[i]`` is the packed input record of full batch ``i``
    (``ids[B*S] | lens[B] | labels[B]``, int32 — the layout of ``B200Engine.d_in``), ``tail`` the ragged last batch."""
    steps: torch.Tensor                                               # [n_full, B*S + 2*B] int32
    tail: Optional[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]]   # (ids [b,S], lens [b], labels [b]) or None

    def __len__(self) -> int:
        return int(self.steps.shape[0]) + (1 if self.tail is not None else 0)


def stage_shuffled_epoch(split: PackedSplit, batch_size: int, generator: Optional[torch.Generator] = None,
                         rank: int = 0, world: int = 1) -> StagedEpoch:
    """Draws the epoch's permutation (the same draw ``iterate_batches(shuffle=True)`` makes) and gathers the WHOLE
    epoch once: one gather per epoch and one small device copy per step instead of a gather and three copies per step
    (the reference re-tokenises every sample every epoch, client1.py:36-45).  The gather is what is staged for the
    next round while a federated aggregation is in flight.

    ``world > 1`` (data-parallel client): every GPU draws the same permutation; global step ``i`` covers
    ``world * batch_size`` consecutive samples of it and GPU ``rank`` takes its ``batch_size`` slice.  The ragged
    remainder of the epoch is then shared out as one smaller last step of equal size per GPU (at most ``world - 1``
    samples are left over), so every GPU runs the same number of steps."""
    n, S = len(split), split.seq_len
    dev = split.input_ids.device
    if generator is not None and generator.device != dev:
        perm = torch.randperm(n, generator=generator).to(dev)
    else:
        perm = torch.randperm(n, generator=generator, device=dev)
    per_step = batch_size * world
    n_full = n // per_step
    ix = perm[: n_full * per_step].view(n_full, world, batch_size)[:, rank, :].reshape(-1)
    steps = torch.empty(n_full, batch_size * S + 2 * batch_size, dtype=torch.int32, device=dev)
    steps[:, : batch_size * S] = split.input_ids[ix].view(n_full, batch_size * S)
    steps[:, batch_size * S: batch_size * S + batch_size] = split.lens[ix].view(n_full, batch_size)
    steps[:, batch_size * S + batch_size:] = split.labels[ix].view(n_full, batch_size)
    tail = None
    rest = perm[n_full * per_step:]
    b = int(rest.numel()) // world
    if b > 0:
        tx = rest[rank * b: (rank + 1) * b]
        tail = (split.input_ids[tx], split.lens[tx], split.labels[tx])
    return StagedEpoch(steps, tail)
