"""Local training and evaluation loops with two interchangeable backends.

Reference: ``train_model`` (client1.py:96-115) and ``evaluate_model`` (client1.py:118-150).  Behaviour kept: Adam
lr 2e-5, batch 16, ``num_epochs`` full passes with a shuffled loader, per-epoch ``Average Loss`` = mean of the
batch-mean losses; evaluation metrics as defined in ``report/metrics.py``.

Backends
* ``B200Backend``  — the CUDA-graph step engine (``models/engine.py``); data stays on the device, the loss is read
  once per epoch instead of twice per step.
* ``TorchBackend`` — the pure-PyTorch module with ``torch.optim.Adam``; used on CPU-only machines (plumbing tier)
  and as the oracle in tests.
"""
from __future__ import annotations

import time
from typing import List, Optional

import torch
import torch.nn as nn

from ..data.pack import PackedSplit, StagedEpoch, iterate_batches, stage_shuffled_epoch
from ..models.arena import ModelDims
from ..ops import native as nv
from ..report.log import EventLog
from ..report.metrics import EvalResult


class TorchBackend:
    name = "torch"

    def __init__(self, dims: ModelDims = ModelDims(), batch: int = 16, seq: int = 128, device: str = "cpu",
                 lr: float = 2e-5, seed: int = 0, init_seed: Optional[int] = 0):
        from ..models.distilbert import DDoSClassifier

        self.dims, self.B, self.S = dims, batch, seq
        self.device = torch.device(device)
        torch.manual_seed(seed)
        self.model = DDoSClassifier(dims, seed=init_seed).to(self.device)
        self.opt = torch.optim.Adam(self.model.parameters(), lr=lr)
        self.crit = nn.CrossEntropyLoss()

    def state_dict(self):
        return self.model.state_dict()

    def load_state_dict(self, sd) -> None:
        self.model.load_state_dict(sd)

    @staticmethod
    def _mask(ids: torch.Tensor, lens: torch.Tensor) -> torch.Tensor:
        return (torch.arange(ids.shape[1], device=ids.device)[None, :] < lens[:, None]).long()

    def stage_epoch(self, split: PackedSplit, generator: Optional[torch.Generator]) -> StagedEpoch:
        return stage_shuffled_epoch(split, self.B, generator)

    def _staged_batches(self, st: StagedEpoch):
        B, S = self.B, self.S
        for rec in st.steps:
            yield rec[: B * S].view(B, S), rec[B * S: B * S + B], rec[B * S + B:]
        if st.tail is not None:
            yield st.tail

    def train_epoch(self, split: PackedSplit, generator: Optional[torch.Generator],
                    staged: Optional[StagedEpoch] = None) -> float:
        self.model.train()
        total, nb = 0.0, 0
        batches = (self._staged_batches(staged) if staged is not None
                   else iterate_batches(split, self.B, shuffle=True, generator=generator))
        for ids, lens, labels in batches:
            ids, lens, labels = ids.to(self.device), lens.to(self.device), labels.to(self.device)
            self.opt.zero_grad()
            loss = self.crit(self.model(ids.long(), self._mask(ids, lens)), labels.long())
            loss.backward()
            self.opt.step()
            total += float(loss.detach())
            nb += 1
        return total / max(nb, 1)

    @torch.no_grad()
    def evaluate(self, split: PackedSplit, want_probs: bool = True) -> EvalResult:
        self.model.eval()
        cm = torch.zeros(2, 2, dtype=torch.int64)
        loss_sum, nb = 0.0, 0
        all_labels: List[int] = []
        all_probs: List[float] = []
        for ids, lens, labels in iterate_batches(split, self.B):
            ids, lens, labels = ids.to(self.device), lens.to(self.device), labels.to(self.device)
            logits = self.model(ids.long(), self._mask(ids, lens))
            loss_sum += float(self.crit(logits, labels.long()))
            nb += 1
            pred = logits.argmax(1)
            for t, p_ in zip(labels.tolist(), pred.tolist()):
                cm[t, p_] += 1
            if want_probs:
                all_labels.extend(labels.tolist())
                all_probs.extend(torch.softmax(logits, 1)[:, 1].tolist())
        return EvalResult.from_confusion(cm.numpy(), loss_sum, nb, all_labels, all_probs)


class B200Backend:
    name = "b200"

    def __init__(self, dims: ModelDims = ModelDims(), batch: int = 16, seq: int = 128, device: int = 0,
                 lr: float = 2e-5, seed: int = 0, init_seed: Optional[int] = 0, flat_params=None, bf16_params=None,
                 capture: bool = True, dp=None):
        from ..models.engine import B200Engine

        self.dims, self.B, self.S = dims, batch, seq
        self.dp = dp
        self.engine = B200Engine(dims, batch, seq, device, lr=lr, seed=seed, init_seed=init_seed,
                                 flat_params=flat_params, bf16_params=bf16_params, dp=dp)
        self.device = self.engine.device
        if capture:
            self.engine.capture()

    def state_dict(self):
        if self.dp is not None:  # collective inside the client group: owners publish their fp32 GEMM weights
            self.dp.allgather_fp32_gemm_params()
        return self.engine.state_dict()

    def load_state_dict(self, sd) -> None:
        self.engine.load_state_dict(sd)

    def stage_epoch(self, split: PackedSplit, generator: Optional[torch.Generator]) -> StagedEpoch:
        """Shuffles and gathers one epoch of ``split`` into per-step input records (device-resident).  Separate from
        ``train_epoch`` so that the next round's first epoch can be staged while a federated aggregation is in
        flight."""
        rank, world = (self.dp.rank, self.dp.world) if self.dp is not None else (0, 1)
        return stage_shuffled_epoch(split, self.B, generator, rank, world)

    def train_epoch(self, split: PackedSplit, generator: Optional[torch.Generator],
                    staged: Optional[StagedEpoch] = None) -> float:
        """One pass over ``split``.  Per step: ONE device copy of the staged input record into the engine's static
        input buffer, then the captured graph — the same work per step as ``bench.py``'s loop.  A data-parallel
        client keeps the ragged end of the epoch (an equal smaller last step on every GPU)."""
        eng = self.engine
        eng.loss_accum.zero_()
        if staged is None:
            staged = self.stage_epoch(split, generator)
        for i in range(staged.steps.shape[0]):
            eng.d_in.copy_(staged.steps[i])
            eng.train_step()
        if staged.tail is not None:  # DataLoader drop_last=False in the reference: a smaller eager step
            self._tail_step(*staged.tail)
        loss, _ = eng.read_loss()
        return loss

    def _tail_step(self, ids, lens, labels) -> None:
        """The last batch of an epoch may be smaller than B (DataLoader drop_last=False in the reference).  The
        captured graph is specialised for B samples, so the tail runs through a lazily-built engine of its own size
        that shares the same parameter/optimizer buffers."""
        from ..models.engine import B200Engine

        b = int(ids.shape[0])
        cache = self.__dict__.setdefault("_tail", {})
        if b not in cache:
            main = self.engine
            e = B200Engine(self.dims, b, self.S, main.device, lr=main.lr, betas=main.betas, eps=main.eps,
                           flat_params=main.arena.flat, bf16_params=main.arena.bf16, init_seed=None, dp=self.dp)
            e.arena.grad, e.arena.exp_avg, e.arena.exp_avg_sq = main.arena.grad, main.arena.exp_avg, main.arena.exp_avg_sq
            e.rng, e.loss_accum = main.rng, main.loss_accum
            e.touched_rows = main.touched_rows  # part of the optimizer state (which vocabulary rows have moments)
            e._bind_params()
            cache[b] = e
        e = cache[b]
        e.set_batch(ids, lens, labels)
        e.train_step()

    def evaluate(self, split: PackedSplit, want_probs: bool = True) -> EvalResult:
        eng = self.engine
        eng.eval_reset()
        all_labels: List[torch.Tensor] = []
        all_probs: List[torch.Tensor] = []
        for ids, lens, labels in iterate_batches(split, self.B):
            b = int(ids.shape[0])
            eng.ids[: b * self.S].copy_(ids.reshape(-1))
            eng.lens[:b].copy_(lens)
            eng.labels[:b].copy_(labels)
            eng.eval_batch(b, want_probs=want_probs)
            if want_probs:
                all_labels.append(labels.clone())
                all_probs.append(eng.probs[:b].clone())
        loss_sum, nb, cm = eng.eval_read()  # the single device->host sync of the evaluation
        labels_l = torch.cat(all_labels).tolist() if all_labels else []
        probs_l = torch.cat(all_probs).tolist() if all_probs else []
        return EvalResult.from_confusion(cm.numpy(), loss_sum, nb, labels_l, probs_l)


def make_backend(prefer: str = "auto", **kw):
    """``auto`` picks the B200 engine whenever a CUDA device and the native library are present."""
    if prefer == "torch" or (prefer == "auto" and not nv.available()):
        kw.pop("flat_params", None)
        kw.pop("bf16_params", None)
        kw.pop("capture", None)
        dev = kw.pop("device", "cpu")
        return TorchBackend(device="cpu" if isinstance(dev, int) and not torch.cuda.is_available() else
                            (f"cuda:{dev}" if isinstance(dev, int) else dev), **kw)
    return B200Backend(**kw)


def stage_first_epoch(backend, train_split: PackedSplit, seed: int):
    """Staging of a round's first epoch ahead of time (while the previous round's aggregation runs): returns what
    ``train_model(..., prestaged=...)`` expects.  Draws from the same seeded generator ``train_model`` would create,
    so the batches are identical to an un-staged run."""
    gen = torch.Generator(device="cpu").manual_seed(seed)
    split = train_split.to(backend.device)
    return gen, backend.stage_epoch(split, gen)


def train_model(backend, train_split: PackedSplit, num_epochs: int = 3, seed: int = 0,
                log: Optional[EventLog] = None, name: str = "Client 1", prestaged=None) -> List[float]:
    """Reference ``train_model``: ``num_epochs`` passes, prints ``Average Loss`` per epoch; returns those losses.
    ``prestaged`` = ``stage_first_epoch(backend, train_split, seed)`` made earlier."""
    log = log or EventLog(name.upper(), quiet=True)
    if prestaged is not None:
        gen, first = prestaged
    else:
        gen, first = torch.Generator(device="cpu").manual_seed(seed), None
    split = train_split.to(backend.device)
    losses = []
    with log.phase("model training", epochs=num_epochs, samples=len(split)):
        for ep in range(num_epochs):
            t0 = time.perf_counter()
            loss = (backend.train_epoch(split, gen, staged=first) if (ep == 0 and first is not None)
                    else backend.train_epoch(split, gen))
            dt = time.perf_counter() - t0
            losses.append(loss)
            print(f"{name} Epoch [{ep + 1}/{num_epochs}], Average Loss: {loss:.4f}", flush=True)
            log.event("epoch", epoch=ep + 1, loss=loss, seconds=dt, samples_per_s=len(split) / dt if dt > 0 else 0.0)
    return losses


def evaluate_model(backend, split: PackedSplit, log: Optional[EventLog] = None, name: str = "Client 1",
                   want_probs: bool = True) -> EvalResult:
    log = log or EventLog(name.upper(), quiet=True)
    with log.phase("model evaluation", samples=len(split)):
        res = backend.evaluate(split.to(backend.device), want_probs=want_probs)
    print(f"{name} Test Accuracy: {res.accuracy:.2f}%, Test Loss: {res.loss:.4f}, Precision: {res.precision:.4f}, "
          f"Recall: {res.recall:.4f}, F1-Score: {res.f1:.4f}", flush=True)
    return res
