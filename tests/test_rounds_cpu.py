"""Deadline + quorum agreement of the NVSwitch path (fed/rounds.py) — run on threads over a HashStore."""
from __future__ import annotations

import threading
import time

import torch.distributed as dist

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.fed.rounds import (
    RoundCoordinator, weights_for)


def _run(n, plan, deadline=0.5, min_clients=None):
    """plan[i] = seconds client i waits before arriving, None = never arrives, 'absent' = announces it sits out"""
    store = dist.HashStore()
    out = [None] * n

    def client(i):
        if plan[i] is None:
            return
        take = plan[i] != "absent"
        time.sleep(0 if not take else plan[i])
        out[i] = RoundCoordinator(store, i, n, deadline_s=deadline, min_clients=min_clients).agree(0, take_part=take)

    ts = [threading.Thread(target=client, args=(i,)) for i in range(n)]
    t0 = time.time()
    [t.start() for t in ts]
    [t.join(10) for t in ts]
    return out, time.time() - t0


def test_everybody_arrives_no_waiting_for_the_deadline():
    out, dt = _run(4, [0, 0.05, 0.1, 0.02], deadline=5.0)
    assert out == [[0, 1, 2, 3]] * 4 and dt < 2.0


def test_dead_client_is_left_out_after_the_deadline_and_all_agree():
    out, dt = _run(4, [0, 0.05, None, 0.02], deadline=0.4, min_clients=2)
    assert out[2] is None and out[0] == out[1] == out[3] == [0, 1, 3] and 0.4 <= dt < 3.0


def test_straggler_after_the_decision_keeps_its_local_model():
    out, _ = _run(3, [0, 0, 1.2], deadline=0.3, min_clients=2)
    assert out[0] == out[1] == [0, 1] and out[2] is None


def test_announced_absence_does_not_cost_the_deadline():
    out, dt = _run(3, [0, "absent", 0.05], deadline=5.0, min_clients=2)
    assert out[0] == out[2] == [0, 2] and out[1] is None and dt < 2.0


def test_quorum_not_met_means_nobody_aggregates():
    out, _ = _run(4, [0, None, None, None], deadline=0.2, min_clients=2)
    assert out == [None] * 4


def test_sample_count_weights():
    assert weights_for([0, 2], None) is None
    assert weights_for([0, 2], [10, 99, 30]) == [0.25, 0.75]
