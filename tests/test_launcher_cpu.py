"""Multi-process launcher on CPU: 2 ranks over gloo (127.0.0.1), tiny model, one FedAvg round."""
from __future__ import annotations

import os
import subprocess
import sys

import pandas as pd
import torch

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.data import synth

PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_rank_gloo_launch(tmp_path):
    tmp = str(tmp_path)
    synth.write_synthetic_csv(os.path.join(tmp, "CICIDS2017.csv"), 200, seed=21)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(29700 + os.getpid() % 200), "-m", f"{PKG}.cli", "launch",
           "--num-clients", "2", "--backend", "torch", "--local-epochs", "1", "--data-path",
           os.path.join(tmp, "CICIDS2017.csv"), "--data-fraction", "0.5", "--out-dir", tmp, "--n-layers", "1", "--dim",
           "64", "--n-heads", "4", "--hidden-dim", "128", "--max-len", "64", "--no-plots", "--quiet"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    a = torch.load(os.path.join(tmp, "client1_model.pth"))
    b = torch.load(os.path.join(tmp, "client2_model.pth"))
    assert all(torch.equal(a[k], b[k]) for k in a)
    assert os.path.exists(os.path.join(tmp, "ddos_distilbert_model.pth"))
    for cid in (1, 2):
        for kind in ("local", "aggregated"):
            assert len(pd.read_csv(os.path.join(tmp, f"client{cid}_{kind}_metrics.csv"))) == 1


def test_clock_sampler_degrades_gracefully_without_a_gpu():
    """bench.py samples clocks through NVML (or nvidia-smi); on a machine with neither it must still return the
    dictionary shape the JSON line expects."""
    from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.utils.clocks import (
        ClockSampler)

    out = ClockSampler(0).start().stop()
    assert set(out) >= {"sm_mhz", "sm_max_mhz", "reasons", "samples"}
    assert isinstance(out["reasons"], list)


def test_three_rank_launch_with_a_client_sitting_out_a_round(tmp_path):
    """Deadline + quorum on the launcher path: client 3 drops out of round 1 (fault plan); clients 1 and 2 aggregate
    without it (quorum 2) and do not wait for a time-out; in round 2 all three average again and end up identical."""
    tmp = str(tmp_path)
    synth.write_synthetic_csv(os.path.join(tmp, "CICIDS2017.csv"), 200, seed=22)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=3", "--master-addr",
           "127.0.0.1", "--master-port", str(29900 + os.getpid() % 100), "-m", f"{PKG}.cli", "launch",
           "--num-clients", "3", "--backend", "torch", "--local-epochs", "1", "--rounds", "2", "--data-path",
           os.path.join(tmp, "CICIDS2017.csv"), "--data-fraction", "0.5", "--out-dir", tmp, "--n-layers", "1", "--dim",
           "64", "--n-heads", "4", "--hidden-dim", "128", "--max-len", "64", "--no-plots", "--quiet",
           "--faults", "drop=3@0", "--min-clients", "2", "--round-timeout", "60"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    assert os.path.exists(os.path.join(tmp, "client1_aggregated_metrics.csv"))
    assert os.path.exists(os.path.join(tmp, "client2_aggregated_metrics.csv"))
    assert not os.path.exists(os.path.join(tmp, "client3_aggregated_metrics.csv"))      # sat round 1 out
    assert os.path.exists(os.path.join(tmp, "client3_aggregated_metrics_round2.csv"))   # back in round 2
    sds = [torch.load(os.path.join(tmp, f"client{c}_model.pth")) for c in (1, 2, 3)]
    assert all(torch.equal(sds[0][k], sds[1][k]) and torch.equal(sds[0][k], sds[2][k]) for k in sds[0])
