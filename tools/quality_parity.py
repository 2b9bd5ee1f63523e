"""BASELINE config 2 at published scale, both implementations on the same data:

    2 clients, a 225,745-row CICIDS2017-shaped CSV (10 % sample per client -> 13,544 train / 4,515 val / 4,515 test
    rows, the published split sizes), batch 16, seq 128, Adam 2e-5, 3 local epochs x 847 steps, one FedAvg round.

  reference : the UNMODIFIED scripts -- `python server.py`, `python client1.py`, `python client2.py` from
              baseline/_ref, three processes in a scratch directory, one GPU per client, their own socket round
              (fp32, no TF32/AMP).  Only stand-ins for matplotlib / seaborn (not installed here) are put on their path.
  ours      : `cli launch --num-clients 2 --transport nvswitch` on the same two GPUs.

Both start from the same random-init DistilBERT encoder (a `./distilbert-base-uncased` directory made once; the
reference reads it with from_pretrained, ours with models/pretrained.py) and the same synthetic vocabulary; shuffling,
dropout streams and the classifier-head init differ, so curves agree within noise, not bit for bit.

    python tools/quality_parity.py prepare  --work W
    python tools/quality_parity.py reference --work W          # needs 2 GPUs (or runs both clients on GPU 0)
    python tools/quality_parity.py ours     --work W
    python tools/quality_parity.py report   --work W --out profiles/config2_quality_parity
"""
from __future__ import annotations

import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline"))
PKG = "detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200"


def prepare(work: str, rows: int) -> None:
    import importlib

    import reference_arm

    os.makedirs(work, exist_ok=True)
    synth = importlib.import_module(f"{PKG}.data.synth")
    csv = os.path.join(work, "CICIDS2017.csv")
    if not os.path.exists(csv):
        synth.write_synthetic_csv(csv, rows, seed=0)
    reference_arm.make_model_dir(os.path.join(work, "distilbert-base-uncased"))
    print("prepared", work)


def run_reference(work: str) -> None:
    import torch

    ref_src = os.path.join(ROOT, "baseline", "_ref")
    cwd = os.path.join(work, "ref")
    os.makedirs(cwd, exist_ok=True)
    for name in ("CICIDS2017.csv", "distilbert-base-uncased"):
        dst = os.path.join(cwd, name)
        if not os.path.exists(dst):
            os.symlink(os.path.join(work, name), dst)
    n_gpu = torch.cuda.device_count()
    env0 = dict(os.environ, PYTHONPATH=os.path.join(ROOT, "baseline", "plot_stubs"), PYTHONUNBUFFERED="1")
    t0 = time.time()
    procs = {}
    logs = {}
    # the reference pickles its CUDA tensors as they are (client1.py:232,284) and the server unpickles them onto the same
    # device index: the server process needs a visible GPU too
    for role, script, gpu in (("server", "server.py", "0" if n_gpu else ""), ("client1", "client1.py", "0"),
                              ("client2", "client2.py", "1" if n_gpu > 1 else "0")):
        env = dict(env0, CUDA_VISIBLE_DEVICES=gpu)
        logs[role] = open(os.path.join(cwd, f"{role}_stdout.txt"), "w")
        procs[role] = subprocess.Popen([sys.executable, os.path.join(ref_src, script)], cwd=cwd, env=env,
                                       stdout=logs[role], stderr=subprocess.STDOUT)
        if role == "server":
            time.sleep(2.0)
    rc = {}
    deadline = time.time() + float(os.environ.get("QUALITY_REF_TIMEOUT", "900"))
    for r, p in procs.items():
        try:
            rc[r] = p.wait(timeout=max(1.0, deadline - time.time()))
        except subprocess.TimeoutExpired:   # e.g. clients polling a server that died: do not burn GPU time on retries
            p.kill()
            rc[r] = "killed at the harness deadline"
    for f in logs.values():
        f.close()
    json.dump({"returncodes": rc, "wall_s": time.time() - t0, "gpus": n_gpu}, open(os.path.join(cwd, "run.json"), "w"))
    print("reference done", rc, f"{time.time() - t0:.1f} s")


def run_ours(work: str, transport: str) -> None:
    out = os.path.join(work, "ours")
    os.makedirs(out, exist_ok=True)
    t0 = time.time()
    cmd = [sys.executable, "-m", f"{PKG}.cli", "launch", "--num-clients", "2", "--transport", transport, "--data-path",
           os.path.join(work, "CICIDS2017.csv"), "--model-dir", os.path.join(work, "distilbert-base-uncased"),
           "--out-dir", out, "--local-epochs", "3", "--event-log", os.path.join(out, "events.jsonl"), "--no-plots"]
    with open(os.path.join(out, "stdout.txt"), "w") as f:
        rc = subprocess.run(cmd, cwd=ROOT, stdout=f, stderr=subprocess.STDOUT, timeout=3600).returncode
    json.dump({"returncode": rc, "wall_s": time.time() - t0, "cmd": cmd}, open(os.path.join(out, "run.json"), "w"))
    print("ours done", rc, f"{time.time() - t0:.1f} s")


_EPOCH = re.compile(r"Epoch \[(\d+)/(\d+)\], Average Loss: ([0-9.]+)")
_TS = re.compile(r"at (\d{4}-\d\d-\d\d \d\d:\d\d:\d\d\.\d+)")


def _csv_row(path):
    import pandas as pd

    return pd.read_csv(path).iloc[0].to_dict() if os.path.exists(path) else None


def _ts(line):
    from datetime import datetime

    m = _TS.search(line)
    return datetime.strptime(m.group(1), "%Y-%m-%d %H:%M:%S.%f").timestamp() if m else None


def report(work: str, out_dir: str) -> None:
    os.makedirs(out_dir, exist_ok=True)
    res = {"reference": {}, "ours": {}}
    ref = os.path.join(work, "ref")
    for cid in (1, 2):
        log = open(os.path.join(ref, f"client{cid}_stdout.txt"), errors="replace").read().splitlines()
        losses = [float(m.group(3)) for ln in log for m in [_EPOCH.search(ln)] if m]
        t_train = [(_ts(ln)) for ln in log if "model training" in ln and _ts(ln)]
        t_send = [(_ts(ln)) for ln in log if ("Starting model send" in ln or "Starting send" in ln) and _ts(ln)]
        t_all = [t for t in map(_ts, log) if t]
        res["reference"][f"client{cid}"] = {
            "epoch_losses": losses,
            "train_s": (t_train[-1] - t_train[0]) if len(t_train) >= 2 else None,
            "whole_run_s": (t_all[-1] - t_all[0]) if t_all else None,
            "local": _csv_row(os.path.join(ref, f"client{cid}_local_metrics.csv")),
            "aggregated": _csv_row(os.path.join(ref, f"client{cid}_aggregated_metrics.csv"))}
        shutil.copy(os.path.join(ref, f"client{cid}_stdout.txt"), os.path.join(out_dir, f"reference_client{cid}_stdout.txt"))
        for kind in ("local", "aggregated"):
            p = os.path.join(ref, f"client{cid}_{kind}_metrics.csv")
            if os.path.exists(p):
                shutil.copy(p, os.path.join(out_dir, f"reference_client{cid}_{kind}_metrics.csv"))
    srv = open(os.path.join(ref, "server_stdout.txt"), errors="replace").read().splitlines()
    ts = [t for t in map(_ts, srv) if t]
    agg = [(_ts(ln)) for ln in srv if "aggregation" in ln.lower() and _ts(ln)]
    res["reference"]["server"] = {"first_to_last_log_s": (ts[-1] - ts[0]) if ts else None,
                                  "aggregate_and_save_s": (agg[-1] - agg[0]) if len(agg) >= 2 else None}
    shutil.copy(os.path.join(ref, "server_stdout.txt"), os.path.join(out_dir, "reference_server_stdout.txt"))
    res["reference"]["run"] = json.load(open(os.path.join(ref, "run.json")))

    ours = os.path.join(work, "ours")
    ev = [json.loads(ln) for ln in open(os.path.join(ours, "events.jsonl"))]
    for cid in (1, 2):
        role = f"CLIENT {cid}"
        mine = [e for e in ev if e.get("role") == role]
        losses = [e["loss"] for e in mine if e.get("event") == "epoch"]
        train = [e for e in mine if e.get("event") == "model training"]
        aggs = [e for e in mine if e.get("event") == "federated aggregation"]
        res["ours"][f"client{cid}"] = {
            "epoch_losses": losses, "train_s": train[0]["seconds"] if train else None,
            "epoch_samples_per_s": [e.get("samples_per_s") for e in mine if e.get("event") == "epoch"],
            "aggregation": {k: aggs[0].get(k) for k in ("seconds", "device_round_ms", "overlapped_host_s", "total_s")}
            if aggs else None,
            "whole_run_s": (mine[-1]["ts"] - mine[0]["ts"]) if mine else None,
            "local": _csv_row(os.path.join(ours, f"client{cid}_local_metrics.csv")),
            "aggregated": _csv_row(os.path.join(ours, f"client{cid}_aggregated_metrics.csv"))}
        for kind in ("local", "aggregated"):
            p = os.path.join(ours, f"client{cid}_{kind}_metrics.csv")
            if os.path.exists(p):
                shutil.copy(p, os.path.join(out_dir, f"ours_client{cid}_{kind}_metrics.csv"))
    res["ours"]["run"] = json.load(open(os.path.join(ours, "run.json")))
    shutil.copy(os.path.join(ours, "events.jsonl"), os.path.join(out_dir, "ours_events.jsonl"))
    shutil.copy(os.path.join(ours, "stdout.txt"), os.path.join(out_dir, "ours_stdout.txt"))
    json.dump(res, open(os.path.join(out_dir, "summary.json"), "w"), indent=1)

    def fmt(v, n=4):
        return "—" if v is None else (f"{v:.{n}f}" if isinstance(v, float) else str(v))

    lines = ["| | reference client 1 | ours client 1 | reference client 2 | ours client 2 |", "|---|---|---|---|---|"]
    cols = [res["reference"]["client1"], res["ours"]["client1"], res["reference"]["client2"], res["ours"]["client2"]]
    for ep in range(3):
        lines.append(f"| train loss, epoch {ep + 1} | " + " | ".join(
            fmt(c["epoch_losses"][ep]) if len(c["epoch_losses"]) > ep else "—" for c in cols) + " |")
    for kind in ("local", "aggregated"):
        for key in ("Accuracy", "Loss", "Precision", "Recall", "F1-Score"):
            lines.append(f"| {kind} test {key} | " + " | ".join(fmt((c[kind] or {}).get(key)) for c in cols) + " |")
    lines.append("| local training wall time (s) | " + " | ".join(fmt(c["train_s"], 1) for c in cols) + " |")
    lines.append("| whole client run (s) | " + " | ".join(fmt(c["whole_run_s"], 1) for c in cols) + " |")
    open(os.path.join(out_dir, "table.md"), "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("cmd", choices=["prepare", "reference", "ours", "report"])
    ap.add_argument("--work", default="/tmp/dcad_quality")
    ap.add_argument("--rows", type=int, default=225745)
    ap.add_argument("--transport", default="nvswitch")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "config2_quality_parity"))
    a = ap.parse_args()
    if a.cmd == "prepare":
        prepare(a.work, a.rows)
    elif a.cmd == "reference":
        run_reference(a.work)
    elif a.cmd == "ours":
        run_ours(a.work, a.transport)
    else:
        report(a.work, a.out)
