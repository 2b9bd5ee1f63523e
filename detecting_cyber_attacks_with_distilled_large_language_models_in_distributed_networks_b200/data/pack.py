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
