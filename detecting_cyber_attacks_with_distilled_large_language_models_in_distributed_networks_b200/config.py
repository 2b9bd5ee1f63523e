"""One configuration object for every role (the reference hard-codes every knob as a literal — SURVEY.md §5.6).

Defaults are the reference's literals: fraction 0.1 (client1.py:23), max_len 128 (:27), batch 16 (:370), lr 2e-5 and
3 epochs (:380), ports 12345/12346 and 2 clients (server.py:11-13), 300 s socket timeout, 5 retries.
"""
from __future__ import annotations

import argparse
import dataclasses
from dataclasses import dataclass, field
from typing import Optional

TRANSPORTS = ("auto", "nvswitch", "nvswitch-oneshot", "nvls", "nvswitch-bf16", "nccl", "socket", "none")


@dataclass
class FedConfig:
    # topology
    num_clients: int = 2
    client_id: int = 1                 # 1-based, like the reference's "Client 1" / "Client 2"
    rounds: int = 1
    transport: str = "socket"          # auto (NVLS from 4 GPUs up, else two-shot) | nvswitch (two-shot) | nvswitch-oneshot |
                                       # nvls | nvswitch-bf16 (bf16 exchange) | nccl (baseline) | socket | none
    gpus_per_client: int = 1
    overlap_rounds: bool = True        # start the aggregation asynchronously and write the local round's files / stage
                                       # the next round's first epoch while it runs (BASELINE config 3)
    # training (reference literals)
    local_epochs: int = 3
    batch_size: int = 16
    max_len: int = 128
    trim_padding: bool = True          # run at the longest tokenised sample's length (rounded up to 32) instead of
                                       # computing columns that are padding in every sample (exact; data/pack.py)
    lr: float = 2e-5
    backend: str = "auto"              # auto | b200 | torch
    seed: Optional[int] = None         # default: 41 + client_id (reference: 42 / 43)
    # data
    data_path: str = "CICIDS2017.csv"
    data_fraction: float = 0.1
    synthetic_rows: int = 0            # >0: generate a synthetic CICIDS-shaped CSV when data_path is missing
    disjoint_shards: bool = False      # extension: clients sample disjoint interleaved shards
    model_dir: str = "./distilbert-base-uncased"   # vocab.txt and model.safetensors / pytorch_model.bin if present
    pretrained: bool = True            # start from the encoder weights in model_dir (random init, logged, if absent)
    # model size overrides (tests / plumbing tier); None = DistilBERT-base
    n_layers: Optional[int] = None
    dim: Optional[int] = None
    n_heads: Optional[int] = None
    hidden_dim: Optional[int] = None
    vocab_size: Optional[int] = None
    # socket transport
    host: str = "localhost"
    port_recv: int = 12345
    port_send: int = 12346
    timeout: float = 300.0
    accept_timeout: Optional[float] = None
    round_timeout: Optional[float] = None  # deadline of a server round's collect / distribute phase (default 5 x timeout)
    allow_remote: bool = False         # socket server: bind a non-loopback --host (the wire protocol is unauthenticated)
    max_retries: int = 5
    gzip_level: int = 0
    min_clients: Optional[int] = None  # quorum for the socket server (default: all)
    faults: Optional[str] = None       # fault-injection plan, e.g. "drop=2;delay=1:0.5;cut=1:1000000@0" (utils/faults.py)
    # outputs
    out_dir: str = "."
    plots: bool = True
    extra_curves: bool = False         # ROC / PR curves (dead code in the reference, optional here)
    save_val_metrics: bool = False     # the reference computes validation metrics and drops them (Q9)
    save_state: bool = False           # also write clientN_state.pt after every round: model + Adam moments + RNG +
                                       # round counter (utils/checkpoint.py); the reference saves the model only
    resume: bool = False               # continue from clientN_state.pt (skips the rounds it already completed)
    event_log: Optional[str] = None
    quiet: bool = False

    def resolved_seed(self) -> int:
        return self.seed if self.seed is not None else 41 + self.client_id

    def model_dims(self):
        from .models.arena import ModelDims

        over = {k: getattr(self, k) for k in ("n_layers", "dim", "n_heads", "hidden_dim", "vocab_size")
                if getattr(self, k) is not None}
        return dataclasses.replace(ModelDims(), **over)


def add_arguments(ap: argparse.ArgumentParser) -> None:
    for f in dataclasses.fields(FedConfig):
        name = "--" + f.name.replace("_", "-")
        if f.type in ("bool", bool):
            ap.add_argument(name, action=argparse.BooleanOptionalAction, default=f.default)
        else:
            base = {"int": int, "float": float, "str": str, "Optional[int]": int, "Optional[float]": float,
                    "Optional[str]": str}.get(str(f.type), str)
            ap.add_argument(name, type=base, default=f.default)


def from_args(ns: argparse.Namespace) -> FedConfig:
    kw = {f.name: getattr(ns, f.name) for f in dataclasses.fields(FedConfig) if hasattr(ns, f.name)}
    cfg = FedConfig(**kw)
    if cfg.transport not in TRANSPORTS:
        raise ValueError(f"transport must be one of {TRANSPORTS}")
    return cfg
