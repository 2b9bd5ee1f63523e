"""Full training-state checkpoints.

The reference only ever saves the model: ``torch.save(model.state_dict(), 'clientN_model.pth')`` (client1.py:388,403;
server.py:77) — no optimizer moments, no RNG state, no epoch / round counter, so a resumed run restarts Adam from zero
and replays different dropout masks.  The interchange files keep that exact format (``fed/client.py``); this module
adds an optional second file that captures everything a bit-faithful resume needs:

    model       the reference-compatible state_dict (102 tensors for DistilBERT-base), on the host
    optimizer   B200 engine: Adam exp_avg / exp_avg_sq as flat fp32 arenas (same layout as the parameter arena) and
                the byte flags of vocabulary rows that have optimizer state;  Torch backend: ``optimizer.state_dict()``
    rng         B200 engine: {seed, step} of the counter-based dropout generator (masks are a pure function of
                them);  Torch backend: the torch CPU (and CUDA) generator states
    progress    round / epoch counters and the DataLoader-shuffle generator state, plus free-form ``extra``

``load_training_state`` restores all of it into a backend built with the same model dimensions.

Data-parallel clients (one client on G GPUs): every GPU holds the Adam moments of the arena ranges it owns only.
``training_state`` / ``load_training_state`` are then COLLECTIVE over the client's group (every rank calls them; the
reporter rank writes the file): the moments are merged into the single-GPU layout on save and split again on load, so
the file does not depend on G.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Any, Dict, Optional

import torch

FORMAT_VERSION = 1


def _host(sd) -> "OrderedDict[str, torch.Tensor]":
    return OrderedDict((k, v.detach().to("cpu").clone()) for k, v in sd.items())


def training_state(backend, round_idx: int = 0, epoch: int = 0, shuffle_generator: Optional[torch.Generator] = None,
                   extra: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
    """Snapshot of ``backend`` (``train.trainer.B200Backend`` or ``TorchBackend``) as plain host objects."""
    st: Dict[str, Any] = {
        "format_version": FORMAT_VERSION,
        "backend": backend.name,
        "model": _host(backend.state_dict()),
        "progress": {"round": int(round_idx), "epoch": int(epoch),
                     "shuffle_generator": shuffle_generator.get_state() if shuffle_generator is not None else None},
        "extra": dict(extra or {}),
    }
    if backend.name == "b200":
        eng = backend.engine
        torch.cuda.synchronize(eng.device)
        exp_avg, exp_avg_sq, touched = eng.gathered_optimizer_state()  # collective inside a data-parallel client
        st["optimizer"] = {"exp_avg": exp_avg.to("cpu"), "exp_avg_sq": exp_avg_sq.to("cpu"),
                           "touched_rows": touched.to("cpu"),
                           "lr": eng.lr, "betas": tuple(eng.betas), "eps": eng.eps}
        seed, step = (int(x) for x in eng.rng.tolist())
        st["rng"] = {"seed": seed, "step": step}
    else:
        st["optimizer"] = backend.opt.state_dict()
        st["rng"] = {"torch_cpu": torch.get_rng_state(),
                     "torch_cuda": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None}
    return st


def save_training_state(path: str, backend, write: bool = True, **kw) -> str:
    """Atomic write (temp file + rename) of ``training_state(backend, **kw)``.  ``write=False``: take part in the
    collectives of a data-parallel client's snapshot without writing the file (every rank but the reporter)."""
    st = training_state(backend, **kw)
    if not write:
        return path
    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save(st, tmp)
    os.replace(tmp, path)
    return path


def load_training_state(path_or_state, backend, shuffle_generator: Optional[torch.Generator] = None) -> Dict[str, Any]:
    """Restore a snapshot into ``backend``; returns its ``progress`` dict (round, epoch) merged with ``extra``."""
    st = path_or_state if isinstance(path_or_state, dict) else torch.load(path_or_state, map_location="cpu",
                                                                          weights_only=False)
    if st.get("format_version") != FORMAT_VERSION:
        raise ValueError(f"unsupported training-state format {st.get('format_version')!r}")
    if st["backend"] != backend.name:
        raise ValueError(f"snapshot was taken from the {st['backend']!r} backend, not {backend.name!r}")
    backend.load_state_dict(st["model"])
    if backend.name == "b200":
        o = st["optimizer"]
        backend.engine.load_optimizer_state(o["exp_avg"], o["exp_avg_sq"], o["touched_rows"], st["rng"]["seed"],
                                            st["rng"]["step"])
    else:
        backend.opt.load_state_dict(st["optimizer"])
        torch.set_rng_state(st["rng"]["torch_cpu"])
        if st["rng"].get("torch_cuda") is not None and torch.cuda.is_available():
            torch.cuda.set_rng_state_all(st["rng"]["torch_cuda"])
    prog = dict(st["progress"])
    if shuffle_generator is not None and prog.get("shuffle_generator") is not None:
        shuffle_generator.set_state(prog["shuffle_generator"])
    prog.update(st.get("extra", {}))
    return prog
