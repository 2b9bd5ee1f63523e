"""A federated client, start to finish — the parametrised replacement for the copy-pasted ``client1.py`` /
``client2.py`` (reference ``main`` client1.py:353-411; call stack in SURVEY.md §3.1).

Per round: train locally -> evaluate (val, test) -> START the aggregation (``transport.aggregate_async``: on the
NVSwitch path the kernel runs on a side stream) and, while it is in flight, write ``clientN_local_metrics.csv`` and
``clientN_model.pth`` and stage the next round's first shuffled epoch -> wait for the aggregation -> evaluate the
aggregated model -> ``clientN_aggregated_metrics.csv`` -> plots -> save again.  (The reference does all of it
strictly one after the other, client1.py:383-403; ``--no-overlap-rounds`` restores that order.)  File names, CSV schema, plot names and the resume-from-``clientN_model.pth`` behaviour match the
reference; with ``rounds > 1`` later rounds add a ``_roundK`` suffix (the reference is single-round).  Degraded paths
are kept: if aggregation fails the client still plots its local results (client1.py:405-410).
"""
from __future__ import annotations

import os
import time
from typing import Dict, Optional, Tuple

import torch

from ..config import FedConfig
from ..data import cicids, synth
from ..data.pack import PackedSplit, executed_seq_len, pack_texts
from ..data.split import split_indices
from ..data.tokenizer import WordPieceTokenizer
from ..report.log import EventLog
from ..report.metrics import EvalResult, save_metrics
from ..report.plots import plot_evaluation
from ..train.trainer import evaluate_model, make_backend, stage_first_epoch, train_model
from .transport import Transport


def prepare_data(cfg: FedConfig, log: EventLog) -> Tuple[PackedSplit, PackedSplit, PackedSplit, WordPieceTokenizer]:
    seed = cfg.resolved_seed()
    path = cfg.data_path
    if not os.path.exists(path):
        if cfg.synthetic_rows <= 0:
            raise FileNotFoundError(f"{path} does not exist (pass --synthetic-rows N to generate CICIDS-shaped data)")
        os.makedirs(cfg.out_dir, exist_ok=True)
        path = os.path.join(cfg.out_dir, "synthetic_CICIDS2017.csv")
        if not os.path.exists(path):
            # several clients of one launch may get here together: each writes its own temporary file (the content is
            # deterministic) and publishes it with an atomic rename, so nobody ever reads a half-written CSV
            tmp = f"{path}.{os.getpid()}.tmp"
            synth.write_synthetic_csv(tmp, cfg.synthetic_rows, seed=0)
            os.replace(tmp, path)
    with log.phase("data preprocessing", path=path):
        shard = (cfg.client_id - 1, cfg.num_clients) if cfg.disjoint_shards else None
        texts, labels = cicids.preprocess_data(path, cfg.data_fraction, seed, shard=shard)
    tok = WordPieceTokenizer.from_dir_or_synthetic(cfg.model_dir)
    with log.phase("tokenisation", samples=len(texts)):
        packed = pack_texts(tok, texts, labels, cfg.max_len)
    tr, va, te = split_indices(len(texts), seed)
    return packed.subset(tr), packed.subset(va), packed.subset(te), tok


def run_client(cfg: FedConfig, transport: Optional[Transport] = None, backend=None,
               backend_kwargs: Optional[dict] = None, reporter: bool = True) -> Dict[str, EvalResult]:
    """``reporter=False``: a non-leading GPU of a data-parallel client — it trains and aggregates (both are collective)
    but leaves evaluation, CSV / PNG / checkpoint writing to the client's first GPU."""
    cid = cfg.client_id
    name = f"Client {cid}"
    os.makedirs(cfg.out_dir, exist_ok=True)
    log = EventLog(f"CLIENT {cid}", cfg.event_log, cfg.quiet)
    log.say("Starting client")
    out = lambda f: os.path.join(cfg.out_dir, f)  # noqa: E731
    model_path = out(f"client{cid}_model.pth")

    train, val, test, tok = prepare_data(cfg, log)
    dims = cfg.model_dims()
    if tok.vocab_size > dims.vocab_size:
        raise ValueError(f"tokenizer vocabulary ({tok.vocab_size}) exceeds the model's ({dims.vocab_size})")
    seq = cfg.max_len
    if cfg.trim_padding:
        # columns that are padding in every sample of every split are dropped (exact: see data/pack.py)
        seq = executed_seq_len(cfg.max_len, max(train.max_len(), val.max_len(), test.max_len()))
        if seq < cfg.max_len:
            log.say(f"longest sample has {max(train.max_len(), val.max_len(), test.max_len())} tokens: running at "
                    f"sequence length {seq} instead of the padded {cfg.max_len}")
            train, val, test = train.trim(seq), val.trim(seq), test.trim(seq)
    if backend is None:
        kw = dict(dims=dims, batch=cfg.batch_size, seq=seq, lr=cfg.lr, seed=cfg.resolved_seed())
        kw.update(backend_kwargs or {})
        backend = make_backend(cfg.backend, **kw)
    if cfg.pretrained and not os.path.exists(model_path):
        # reference client1.py:55-56: the encoder starts from the pretrained DistilBERT in model_dir (when present)
        from ..models.pretrained import load_pretrained_encoder

        load_pretrained_encoder(backend, cfg.model_dir, log)
    if os.path.exists(model_path):  # reference client1.py:375-377: continue from the last (aggregated) weights
        log.say(f"loading pre-trained model from {model_path}")
        backend.load_state_dict(torch.load(model_path, map_location="cpu"))

    state_path = out(f"client{cid}_state.pt")
    first_round = 0
    if cfg.resume and os.path.exists(state_path):
        from ..utils.checkpoint import load_training_state

        prog = load_training_state(state_path, backend)
        first_round = int(prog.get("round", 0))
        log.say(f"resumed full training state from {state_path}: {first_round} round(s) already done")

    results: Dict[str, EvalResult] = {}
    prestaged = None
    for rnd in range(first_round, cfg.rounds):
        sfx = "" if rnd == 0 else f"_round{rnd + 1}"
        train_model(backend, train, cfg.local_epochs, seed=cfg.resolved_seed() + 1000 * rnd, log=log, name=name,
                    prestaged=prestaged)
        prestaged = None
        sd = backend.state_dict()  # (collective inside a data-parallel client group); a snapshot, not a view
        local = val_local = None
        if reporter:
            print(f"{name} evaluating local model on validation set...")
            val_local = evaluate_model(backend, val, log, name, want_probs=False)
            print(f"{name} evaluating local model on test set...")
            local = evaluate_model(backend, test, log, name, want_probs=cfg.extra_curves)
            results[f"local{sfx}"] = local

        def write_local_files() -> None:
            if not reporter:
                return
            save_metrics(local, out(f"client{cid}_local_metrics{sfx}.csv"))
            if cfg.save_val_metrics:
                save_metrics(val_local, out(f"client{cid}_local_val_metrics{sfx}.csv"))
            torch.save(sd, model_path)

        aggregated: Optional[EvalResult] = None
        ok = False
        if transport is not None and cfg.overlap_rounds:
            with log.phase("federated aggregation", transport=transport.name, round=rnd + 1, overlapped=True) as ex:
                t0 = time.perf_counter()
                pending = transport.aggregate_async(backend)   # NVSwitch: the kernel is now running on a side stream
                ex["launched_after_s"] = time.perf_counter() - t0
                # ---- work that does not depend on the aggregated model, done while the GPUs average ----
                write_local_files()
                if rnd + 1 < cfg.rounds and hasattr(backend, "stage_epoch"):
                    prestaged = stage_first_epoch(backend, train, cfg.resolved_seed() + 1000 * (rnd + 1))
                ex["overlapped_host_s"] = time.perf_counter() - t0
                ok = pending.wait()
                ex["total_s"] = time.perf_counter() - t0
                if getattr(transport, "last_round_ms", None) is not None:
                    ex["device_round_ms"] = transport.last_round_ms
        else:
            write_local_files()
            if transport is not None:
                with log.phase("federated aggregation", transport=transport.name, round=rnd + 1):
                    ok = transport.aggregate(backend)
        del sd
        if ok and not reporter:
            backend.state_dict()  # keep the collective call pattern of the reporter
        if ok and reporter:
            log.say("updated with aggregated model")
            print(f"{name} evaluating aggregated model on validation set...")
            val_agg = evaluate_model(backend, val, log, name, want_probs=False)
            print(f"{name} evaluating aggregated model on test set...")
            aggregated = evaluate_model(backend, test, log, name, want_probs=cfg.extra_curves)
            save_metrics(aggregated, out(f"client{cid}_aggregated_metrics{sfx}.csv"))
            if cfg.save_val_metrics:
                save_metrics(val_agg, out(f"client{cid}_aggregated_val_metrics{sfx}.csv"))
            torch.save(backend.state_dict(), model_path)
            results[f"aggregated{sfx}"] = aggregated
        elif transport is not None:
            log.say("skipping aggregated model evaluation due to aggregation failure")
        if cfg.plots and reporter:
            with log.phase("plot generation"):
                plot_evaluation(local, aggregated, out(f"client{cid}_plots{sfx}"), cid, extra_curves=cfg.extra_curves)
        if cfg.save_state and (reporter or getattr(backend, "dp", None) is not None):
            from ..utils.checkpoint import save_training_state

            # data-parallel client: every GPU of the group joins the merge of the sharded Adam moments, one writes
            save_training_state(state_path, backend, write=reporter, round_idx=rnd + 1, epoch=cfg.local_epochs)
    log.say("Client shutdown")
    return results
