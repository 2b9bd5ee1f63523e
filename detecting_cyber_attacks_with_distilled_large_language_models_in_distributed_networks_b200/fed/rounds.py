"""Who takes part in a federated round: a deadline + quorum agreement over a key-value store.

The reference's server waits for exactly ``NUM_CLIENTS`` uploads and refuses to aggregate otherwise
(server.py:69-71,124-132); a client that dies leaves everybody blocked until a 300 s socket timeout.  On the NVSwitch
path there is no server process to decide, and the aggregation kernels of all participants must be launched with the
SAME participant list (they split the arena into one shard per participant and handshake with each other), so the
clients agree among themselves:

1. a client that has finished local training marks itself ready for round ``r``;
2. every client polls the ready marks; whoever first sees all clients ready — or its own deadline pass — proposes
   the set it has seen with an atomic compare-and-set on the round's decision key; the first proposal wins and
   everybody (late arrivals included) reads that one decision;
3. a client that is not in the decided set keeps its local model for this round (the reference's degraded path,
   client1.py:405-410) and rejoins at the next round; a decision with fewer than ``min_clients`` members means
   nobody aggregates.

The store is ``torch.distributed``'s rendezvous store (a TCPStore owned by the launcher, so it outlives any one
client) in production and a ``HashStore`` in the CPU tests.  Only a few bytes per round travel through it.
"""
from __future__ import annotations

import time
from typing import List, Optional, Sequence


class RoundCoordinator:
    def __init__(self, store, client_idx: int, num_clients: int, deadline_s: float = 300.0,
                 min_clients: Optional[int] = None, poll_s: float = 0.01, prefix: str = "fed"):
        self.store, self.idx, self.n = store, client_idx, num_clients
        self.deadline_s = deadline_s
        self.min_clients = num_clients if min_clients is None else max(1, min_clients)
        self.poll_s, self.prefix = poll_s, prefix
        self.last_wait_s = 0.0

    def _key(self, rnd: int, what: str) -> str:
        return f"{self.prefix}/r{rnd}/{what}"

    def _ready_now(self, rnd: int) -> List[int]:
        return [j for j in range(self.n) if self.store.check([self._key(rnd, f"ready/{j}")])]

    def agree(self, rnd: int, take_part: bool = True) -> Optional[List[int]]:
        """Blocks until round ``rnd``'s participant set is decided; returns it (sorted client indices) when this
        client is a member and the quorum is met, else ``None``.  ``take_part=False``: this client sits the round out
        (fault injection, a client that failed locally) but still lets the others decide without waiting for it."""
        t0 = time.time()
        if take_part:
            self.store.set(self._key(rnd, f"ready/{self.idx}"), b"1")
        else:
            self.store.set(self._key(rnd, f"absent/{self.idx}"), b"1")
        dkey = self._key(rnd, "decision")
        decided: Optional[bytes] = None
        while decided is None:
            if self.store.check([dkey]):
                decided = self.store.get(dkey)
                break
            ready = self._ready_now(rnd)
            absent = [j for j in range(self.n) if self.store.check([self._key(rnd, f"absent/{j}")])]
            everyone_heard = len(ready) + len(absent) >= self.n
            if everyone_heard or time.time() - t0 >= self.deadline_s:
                proposal = ",".join(str(j) for j in sorted(ready)).encode() or b"-"
                decided = self.store.compare_set(dkey, "", proposal)  # first proposal wins; returns the winner
                break
            time.sleep(self.poll_s)
        self.last_wait_s = time.time() - t0
        members = [] if decided in (b"-", b"") else [int(x) for x in decided.decode().split(",")]
        if not take_part or self.idx not in members or len(members) < self.min_clients:
            return None
        return members


def weights_for(members: Sequence[int], sample_counts: Optional[Sequence[int]]) -> Optional[List[float]]:
    """FedAvg sample-count weights n_k / n restricted to the participating clients (None: the reference's unweighted
    mean, server.py:73-76)."""
    if sample_counts is None:
        return None
    tot = float(sum(sample_counts[m] for m in members))
    return [sample_counts[m] / tot for m in members]
