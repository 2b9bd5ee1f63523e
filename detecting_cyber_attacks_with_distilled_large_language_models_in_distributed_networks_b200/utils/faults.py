"""Fault injection for the federated round (the reference has none; its only defences are 300 s socket
timeouts, 5 download attempts and a 5-failure send budget — client1.py:22,314-336, server.py:92-112).

A ``FaultPlan`` describes what should go wrong and for whom; transports ask it at three points:

* ``before_upload(client, round)``   -> False: the client silently never uploads this round (crashed / partitioned)
* ``delay(client, round)``           -> seconds to sleep before the upload (straggler)
* ``cut_after(client, round)``       -> byte count after which the upload connection is torn down (killed mid-transfer)

The plan is plain data so it can come from a test, from ``FedConfig.faults`` or from the environment
(``DCAD_FAULTS``), using one small grammar::

    drop=2            client 2 never uploads
    drop=2@0,1        ... in rounds 0 and 1 only
    delay=1:0.5       client 1 sleeps 0.5 s before uploading
    cut=1:1000000     client 1's upload is aborted after 1 000 000 payload bytes
    cut=1:1000000@0   ... in round 0 only

Clauses are separated by ``;``.  Client ids are the 1-based ids of ``FedConfig.client_id``.
"""
from __future__ import annotations

import os
import time
from dataclasses import dataclass, field
from typing import Dict, Optional, Set, Tuple


class InjectedFault(ConnectionError):
    """Raised by a transport when the plan tells it to break a transfer."""


_Rounds = Optional[Set[int]]  # None = every round


@dataclass
class FaultPlan:
    drop: Dict[int, _Rounds] = field(default_factory=dict)
    delays: Dict[int, Tuple[float, _Rounds]] = field(default_factory=dict)
    cuts: Dict[int, Tuple[int, _Rounds]] = field(default_factory=dict)
    fired: list = field(default_factory=list)  # (kind, client, round) log, for tests and the event log

    # ------------------------------------------------------------------ construction
    @staticmethod
    def parse(spec: Optional[str]) -> "FaultPlan":
        plan = FaultPlan()
        if not spec:
            return plan
        for clause in spec.split(";"):
            clause = clause.strip()
            if not clause:
                continue
            kind, _, rest = clause.partition("=")
            body, _, rounds_s = rest.partition("@")
            rounds: _Rounds = {int(r) for r in rounds_s.split(",")} if rounds_s else None
            kind = kind.strip().lower()
            if kind == "drop":
                for c in body.split(","):
                    plan.drop[int(c)] = rounds
            elif kind == "delay":
                c, _, secs = body.partition(":")
                plan.delays[int(c)] = (float(secs), rounds)
            elif kind == "cut":
                c, _, nbytes = body.partition(":")
                plan.cuts[int(c)] = (int(nbytes), rounds)
            else:
                raise ValueError(f"unknown fault clause {clause!r} (drop= | delay= | cut=)")
        return plan

    @staticmethod
    def from_env(var: str = "DCAD_FAULTS") -> "FaultPlan":
        return FaultPlan.parse(os.environ.get(var))

    def __bool__(self) -> bool:
        return bool(self.drop or self.delays or self.cuts)

    # ------------------------------------------------------------------ queries
    @staticmethod
    def _hit(rounds: _Rounds, rnd: int) -> bool:
        return rounds is None or rnd in rounds

    def before_upload(self, client: int, rnd: int = 0) -> bool:
        """False when the client must not upload at all this round."""
        if client in self.drop and self._hit(self.drop[client], rnd):
            self.fired.append(("drop", client, rnd))
            return False
        return True

    def delay(self, client: int, rnd: int = 0, sleep: bool = True) -> float:
        secs = 0.0
        if client in self.delays and self._hit(self.delays[client][1], rnd):
            secs = self.delays[client][0]
            self.fired.append(("delay", client, rnd))
            if sleep and secs > 0:
                time.sleep(secs)
        return secs

    def cut_after(self, client: int, rnd: int = 0) -> Optional[int]:
        if client in self.cuts and self._hit(self.cuts[client][1], rnd):
            self.fired.append(("cut", client, rnd))
            return self.cuts[client][0]
        return None
