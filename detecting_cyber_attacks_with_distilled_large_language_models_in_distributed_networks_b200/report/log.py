"""Structured event log (JSONL) + wall-clock phase prints.

The reference's only observability is ``print(f"[CLIENT 1] Starting ... at {datetime.now()}")`` around every phase
(client1.py:85,92,97,115,...; server.py:30,44,...).  The same phase lines are kept on stdout (they are what a user of
the reference greps for) and every phase is also appended to a JSONL file with its duration, optional device time,
throughput and payload sizes — the fields BASELINE.json's metrics need.  NVTX ranges are emitted when CUDA is present
so the phases show up in profiler timelines.
"""
from __future__ import annotations

import contextlib
import json
import os
import time
from datetime import datetime
from typing import Any, Dict, Optional


class EventLog:
    def __init__(self, role: str, path: Optional[str] = None, quiet: bool = False):
        self.role, self.path, self.quiet = role, path, quiet
        if path:
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)

    def say(self, msg: str) -> None:
        if not self.quiet:
            print(f"[{self.role}] {msg} at {datetime.now()}", flush=True)

    def event(self, name: str, **fields: Any) -> None:
        rec: Dict[str, Any] = {"ts": time.time(), "role": self.role, "event": name}
        rec.update(fields)
        if self.path:
            with open(self.path, "a") as f:
                f.write(json.dumps(rec, default=float) + "\n")

    @contextlib.contextmanager
    def phase(self, name: str, **fields: Any):
        self.say(f"Starting {name}")
        nvtx = None
        try:
            import torch

            if torch.cuda.is_available():
                torch.cuda.nvtx.range_push(f"{self.role}:{name}")
                nvtx = torch.cuda.nvtx
        except Exception:  # noqa: BLE001
            nvtx = None
        t0 = time.perf_counter()
        extra: Dict[str, Any] = {}
        try:
            yield extra
        finally:
            dt = time.perf_counter() - t0
            if nvtx is not None:
                nvtx.range_pop()
            self.event(name, seconds=dt, **fields, **extra)
            self.say(f"Finished {name} ({dt:.3f} s)")
