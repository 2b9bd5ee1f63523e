"""Fault injection (utils/faults.py) against the socket server's deadline + quorum and the loopback transport:
a client that drops out, a straggler, and an upload killed mid-transfer."""
from __future__ import annotations

import socket
import threading
from collections import OrderedDict

import torch

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.fed.transport import (
    LoopbackTransport)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel import (
    socket_transport as st)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.report.log import EventLog
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.utils.faults import (
    FaultPlan)


def _free_port():
    s = socket.socket()
    s.bind(("localhost", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _sd(seed, n=3, rows=64):
    g = torch.Generator().manual_seed(seed)
    return OrderedDict((f"w{i}", torch.randn(rows, 33, generator=g)) for i in range(n))


def test_plan_grammar():
    plan = FaultPlan.parse("drop=2@0,1; delay=1:0.25; cut=3:1000@2")
    assert not plan.before_upload(2, 0) and not plan.before_upload(2, 1) and plan.before_upload(2, 2)
    assert plan.before_upload(1, 0)
    assert plan.delay(1, 5, sleep=False) == 0.25 and plan.delay(2, 0, sleep=False) == 0.0
    assert plan.cut_after(3, 2) == 1000 and plan.cut_after(3, 0) is None
    assert ("drop", 2, 0) in plan.fired and ("cut", 3, 2) in plan.fired
    assert not FaultPlan.parse(None) and not FaultPlan.parse("")


def test_server_survives_cut_upload_and_dropout_with_quorum():
    """Client 2's upload dies mid-transfer and client 3 never shows up; with a deadline and a quorum of 1 the server
    aggregates what arrived (client 1's model) and serves it."""
    pr, ps = _free_port(), _free_port()
    srv = st.FedServer(3, port_recv=pr, port_send=ps, save_path=None, log=EventLog("SERVER", quiet=True),
                       accept_timeout=1.5, min_clients=1, timeout=5)
    srv.open()
    out = {}
    t = threading.Thread(target=lambda: out.setdefault("agg", srv.run_round()), daemon=True)
    t.start()
    plan = FaultPlan.parse("cut=2:4096; drop=3")
    sds = {i: _sd(i, rows=512) for i in (1, 2, 3)}
    res = {}

    def client(i):
        ok = st.send_model(sds[i], "localhost", pr, timeout=10, faults=plan, client_id=i)
        res[i] = ok
        if ok:
            res[f"agg{i}"] = st.receive_aggregated_model("localhost", ps, timeout=10, max_retries=2)

    cs = [threading.Thread(target=client, args=(i,)) for i in (1, 2, 3)]
    [c.start() for c in cs]
    [c.join(30) for c in cs]
    t.join(30)
    srv.close()
    assert res[1] is True and res[2] is False and res[3] is False
    assert {k for k, *_ in plan.fired} == {"cut", "drop"}
    agg = res["agg1"]
    assert agg is not None and all(torch.equal(agg[k], sds[1][k]) for k in sds[1])


def test_server_refuses_below_quorum():
    pr, ps = _free_port(), _free_port()
    srv = st.FedServer(2, port_recv=pr, port_send=ps, save_path=None, log=EventLog("SERVER", quiet=True),
                       accept_timeout=0.8, timeout=5)        # quorum defaults to all clients
    srv.open()
    out = {}
    t = threading.Thread(target=lambda: out.setdefault("agg", srv.run_round()), daemon=True)
    t.start()
    plan = FaultPlan.parse("drop=2")
    ok1 = st.send_model(_sd(1), "localhost", pr, timeout=10, faults=plan, client_id=1)
    ok2 = st.send_model(_sd(2), "localhost", pr, timeout=10, faults=plan, client_id=2)
    t.join(30)
    srv.close()
    assert ok1 and not ok2 and out["agg"] is None


class _Holder:
    def __init__(self, sd):
        self.sd = sd

    def state_dict(self):
        return self.sd

    def load_state_dict(self, sd):
        self.sd = OrderedDict((k, v.clone()) for k, v in sd.items())


def test_loopback_quorum_round():
    plan = FaultPlan.parse("drop=3@0")
    tr = LoopbackTransport(3, faults=plan, min_clients=2)
    hs = {i: _Holder(_sd(i)) for i in (1, 2, 3)}
    orig = {i: OrderedDict((k, v.clone()) for k, v in hs[i].sd.items()) for i in hs}
    assert tr.aggregate(hs[1], client_id=1) and tr.aggregate(hs[2], client_id=2)
    assert tr.aggregate(hs[3], client_id=3) is False          # dropped by the plan: never registers
    assert tr.finish_round() is True                          # deadline: 2 of 3 meets the quorum
    for k in orig[1]:
        exp = (orig[1][k] + orig[2][k]) / 2
        assert torch.allclose(hs[1].sd[k], exp) and torch.allclose(hs[2].sd[k], exp)
        assert torch.equal(hs[3].sd[k], orig[3][k])           # the dropout keeps its local model
    # next round: nobody is dropped, the third arrival closes the round by itself
    assert tr.aggregate(hs[1], client_id=1) and tr.aggregate(hs[2], client_id=2) and tr.aggregate(hs[3], client_id=3)
    assert all(torch.equal(hs[1].sd[k], hs[3].sd[k]) for k in orig[1])
