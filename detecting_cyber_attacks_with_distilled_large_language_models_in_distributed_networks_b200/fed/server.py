"""Socket FedAvg server entry point (reference ``server.py`` main :116-137), multi-round and quorum capable."""
from __future__ import annotations

import os

from ..config import FedConfig
from ..parallel.socket_transport import FedServer
from ..report.log import EventLog


def run_server(cfg: FedConfig) -> None:
    os.makedirs(cfg.out_dir, exist_ok=True)
    log = EventLog("SERVER", cfg.event_log, cfg.quiet)
    srv = FedServer(num_clients=cfg.num_clients, host=cfg.host, port_recv=cfg.port_recv, port_send=cfg.port_send,
                    timeout=cfg.timeout, accept_timeout=cfg.accept_timeout, gzip_level=cfg.gzip_level,
                    save_path=os.path.join(cfg.out_dir, "ddos_distilbert_model.pth"), min_clients=cfg.min_clients,
                    log=log, round_timeout=cfg.round_timeout, allow_remote=cfg.allow_remote)
    srv.serve(cfg.rounds)
