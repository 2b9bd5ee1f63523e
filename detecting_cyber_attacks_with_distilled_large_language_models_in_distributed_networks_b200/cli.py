"""Command line: ``python -m <package>.cli {client,server,launch,synth-data,make-vocab}``.

    server      socket FedAvg server (cross-host fallback; wire-compatible with the reference's clients)
    client      one federated client over the socket transport (wire-compatible with the reference's server)
    launch      N clients on N GPUs of this box, aggregation over NVSwitch peer memory (or --transport nccl baseline)
    synth-data  write a synthetic CICIDS2017-shaped CSV
    make-vocab  write the synthetic 30,522-entry WordPiece vocab.txt
"""
from __future__ import annotations

import argparse
import sys

from . import config as C


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="ddosfed")
    sub = ap.add_subparsers(dest="cmd", required=True)
    for name in ("client", "server", "launch"):
        C.add_arguments(sub.add_parser(name))
    p = sub.add_parser("synth-data")
    p.add_argument("--out", default="CICIDS2017.csv")
    p.add_argument("--rows", type=int, default=225745)
    p.add_argument("--seed", type=int, default=0)
    p = sub.add_parser("make-vocab")
    p.add_argument("--out", default="distilbert-base-uncased/vocab.txt")
    ns = ap.parse_args(argv)

    if ns.cmd == "synth-data":
        from .data.synth import write_synthetic_csv

        print(write_synthetic_csv(ns.out, ns.rows, seed=ns.seed))
        return 0
    if ns.cmd == "make-vocab":
        from .data.tokenizer import write_vocab

        print(write_vocab(ns.out))
        return 0
    cfg = C.from_args(ns)
    if ns.cmd == "server":
        from .fed.server import run_server

        run_server(cfg)
    elif ns.cmd == "client":
        from .fed.client import run_client
        from .fed.transport import SocketTransport
        from .report.log import EventLog

        tr = None
        if cfg.transport == "socket":
            from .utils.faults import FaultPlan

            plan = FaultPlan.parse(cfg.faults) if cfg.faults else FaultPlan.from_env()
            tr = SocketTransport(cfg.host, cfg.port_recv, cfg.port_send, cfg.timeout, cfg.max_retries, cfg.gzip_level,
                                 EventLog(f"CLIENT {cfg.client_id}", cfg.event_log, cfg.quiet), faults=plan,
                                 client_id=cfg.client_id)
        run_client(cfg, tr)
    elif ns.cmd == "launch":
        from .fed.launcher import launch

        launch(cfg)
    return 0


if __name__ == "__main__":
    sys.exit(main())
