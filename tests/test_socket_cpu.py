"""Socket fallback transport: framing, our client <-> our server, and interop with the reference's code in both
directions over loopback."""
from __future__ import annotations

import socket
import threading
import time
from collections import OrderedDict

import pytest
import torch

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel import (
    socket_transport as st)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.report.log import EventLog


def _free_port():
    s = socket.socket()
    s.bind(("localhost", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _sd(seed, n=3):
    g = torch.Generator().manual_seed(seed)
    return OrderedDict((f"w{i}", torch.randn(64, 33, generator=g)) for i in range(n))


def test_compress_roundtrip_and_reference_interop(ref_server):
    sd = _sd(0)
    blob = st.compress_data(sd, level=1)
    back = ref_server.decompress_data(blob)            # the reference can read what we write
    assert all(torch.equal(back[k], sd[k]) for k in sd)
    back2 = st.decompress_data(ref_server.compress_data(sd))  # and we can read what it writes
    assert all(torch.equal(back2[k], sd[k]) for k in sd)


def test_two_clients_one_round_and_probe_tolerance():
    pr, ps = _free_port(), _free_port()
    srv = st.FedServer(2, port_recv=pr, port_send=ps, save_path=None, log=EventLog("SERVER", quiet=True),
                       aggregate=None)
    srv.open()
    t = threading.Thread(target=srv.run_round, daemon=True)
    t.start()
    # a reference-style readiness probe (connect + close) must not consume a client slot
    for port in (pr, ps):
        probe = socket.create_connection(("localhost", port), timeout=3)
        probe.close()
    results = {}

    def client(i):
        ok = st.send_model(_sd(i), "localhost", pr, timeout=30)
        agg = st.receive_aggregated_model("localhost", ps, timeout=30) if ok else None
        results[i] = (ok, agg)

    cs = [threading.Thread(target=client, args=(i,)) for i in (1, 2)]
    [c.start() for c in cs]
    [c.join(60) for c in cs]
    t.join(60)
    srv.close()
    a, b = _sd(1), _sd(2)
    for i in (1, 2):
        ok, agg = results[i]
        assert ok and agg is not None
        for k in a:
            # upload order is a race, and fp32 addition of two numbers commutes, so either order is bit-identical
            assert torch.equal(agg[k], (a[k] + b[k]) / 2)


def test_our_client_against_reference_server(ref_server, tmp_path, monkeypatch):
    """Wire interop: the reference's handle_client_recv / send_chunked_data serve a client from this package."""
    monkeypatch.chdir(tmp_path)
    pr, ps = _free_port(), _free_port()
    ref_server.received_models.clear()
    ref_server.NUM_CLIENTS = 1
    out = {}

    def serve():
        ls = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        ls.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        ls.bind(("localhost", pr))
        ls.listen(1)
        conn, addr = ls.accept()
        ref_server.handle_client_recv(conn, addr, 0)       # reference upload path
        ls.close()
        agg = ref_server.aggregate_models()                 # reference FedAvg (N=1 -> identity)
        ds = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        ds.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        ds.bind(("localhost", ps))
        ds.listen(1)
        c2, _ = ds.accept()
        out["ack"] = ref_server.send_chunked_data(c2, ref_server.compress_data(agg))  # reference download path
        c2.close()
        ds.close()

    t = threading.Thread(target=serve, daemon=True)
    t.start()
    sd = _sd(5)
    assert st.send_model(sd, "localhost", pr, timeout=30)
    agg = st.receive_aggregated_model("localhost", ps, timeout=30)
    t.join(30)
    assert out.get("ack") is True and agg is not None
    assert all(torch.equal(agg[k], sd[k]) for k in sd)
    ref_server.received_models.clear()


def test_reference_client_functions_against_our_server(ref_client1):
    """The reference's send_chunked_data / receive_chunked_data / (de)compress talk to FedServer."""
    pr, ps = _free_port(), _free_port()
    srv = st.FedServer(1, port_recv=pr, port_send=ps, save_path=None, log=EventLog("SERVER", quiet=True))
    srv.open()
    t = threading.Thread(target=srv.run_round, daemon=True)
    t.start()
    sd = _sd(9)
    s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    s.connect(("localhost", pr))
    assert ref_client1.send_chunked_data(s, ref_client1.compress_data(sd)) is True
    s.close()
    s2 = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
    s2.settimeout(30)
    s2.connect(("localhost", ps))
    data = ref_client1.receive_chunked_data(s2)
    s2.close()
    t.join(30)
    srv.close()
    agg = ref_client1.decompress_data(data)
    assert all(torch.equal(agg[k], sd[k]) for k in sd)


def test_send_model_reports_failure_when_no_server():
    assert st.send_model(_sd(1), "localhost", _free_port(), timeout=0.5) is False
    assert st.receive_aggregated_model("localhost", _free_port(), max_retries=1, timeout=0.5) is None


def test_quorum_round_with_missing_client():
    pr, ps = _free_port(), _free_port()
    srv = st.FedServer(2, port_recv=pr, port_send=ps, save_path=None, accept_timeout=1.5, min_clients=1,
                       log=EventLog("SERVER", quiet=True))
    srv.open()
    t = threading.Thread(target=srv.run_round, daemon=True)
    t.start()
    sd = _sd(3)
    assert st.send_model(sd, "localhost", pr, timeout=10)
    agg = st.receive_aggregated_model("localhost", ps, timeout=10)
    t.join(30)
    srv.close()
    assert agg is not None and all(torch.equal(agg[k], sd[k]) for k in sd)


def test_untrusted_payloads_are_refused():
    """The wire decoder accepts a pickled state_dict and nothing else: a payload that names any other global (the
    classic ``os.system`` reduce) is rejected before anything runs, also when it hides inside a tensor's storage."""
    import gzip
    import io
    import os
    import pickle

    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel import (
        socket_transport as st)

    class Evil:
        def __reduce__(self):
            return (os.system, ("echo pwned > /tmp/dcad_pwned",))

    def gz(b):
        buf = io.BytesIO()
        with gzip.GzipFile(fileobj=buf, mode="wb", compresslevel=0) as f:
            f.write(b)
        return buf.getvalue()

    if os.path.exists("/tmp/dcad_pwned"):
        os.remove("/tmp/dcad_pwned")
    with pytest.raises(pickle.UnpicklingError):
        st.decompress_data(gz(pickle.dumps(Evil())))
    # nested: a "storage" whose bytes are themselves a malicious pickle (what torch.storage._load_from_bytes would run)
    nested = pickle.dumps(Evil())

    class FakeStorage:
        def __reduce__(self):
            import torch.storage
            return (torch.storage._load_from_bytes, (nested,))

    with pytest.raises(Exception):
        st.decompress_data(gz(pickle.dumps(FakeStorage())))
    assert not os.path.exists("/tmp/dcad_pwned")
    # the legitimate payload (tensors of several dtypes, a Parameter, the OrderedDict's _metadata) still round-trips
    from collections import OrderedDict
    sd = OrderedDict(a=torch.randn(3, 4), b=torch.arange(5), c=torch.randn(2).bfloat16(),
                     p=torch.nn.Parameter(torch.randn(2)))
    sd._metadata = OrderedDict({"": {"version": 1}})
    back = st.decompress_data(st.compress_data(sd))
    assert list(back) == list(sd) and back._metadata == sd._metadata
    assert all(torch.equal(back[k], sd[k]) for k in sd)


def test_server_refuses_remote_bind_and_oversized_payloads():
    import socket as pysocket

    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel import (
        socket_transport as st)

    with pytest.raises(ValueError):
        st.FedServer(num_clients=1, host="0.0.0.0", save_path=None)
    st.FedServer(num_clients=1, host="0.0.0.0", save_path=None, allow_remote=True)  # explicit opt-in is accepted
    a, b = pysocket.socketpair()
    try:
        a.sendall(b"99999999999\n")
        with pytest.raises(ValueError):
            st.receive_chunked_data(b)
    finally:
        a.close()
        b.close()


def test_round_deadline_applies_quorum(tmp_path):
    """A multi-round server whose second client never shows up: the collect phase ends at the round deadline, the
    quorum (min_clients=1) is applied and the round completes with the client that did come."""
    import threading

    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel import (
        socket_transport as st)

    pr, ps = _free_port(), _free_port()
    srv = st.FedServer(num_clients=2, port_recv=pr, port_send=ps, timeout=5, round_timeout=1.5, min_clients=1,
                       save_path=None, aggregate=lambda sds: sds[0])
    srv.open()
    out = {}
    t = threading.Thread(target=lambda: out.setdefault("agg", srv.run_round()), daemon=True)
    t.start()
    sd = {"w": torch.ones(4)}
    assert st.send_model(sd, port=pr, timeout=5)
    got = st.receive_aggregated_model(port=ps, timeout=10, max_retries=2)
    t.join(timeout=20)
    srv.close()
    assert not t.is_alive() and got is not None and torch.equal(got["w"], sd["w"])
