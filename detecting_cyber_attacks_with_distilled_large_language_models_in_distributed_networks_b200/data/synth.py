"""Synthetic CICIDS2017-shaped data for boxes without the real capture (no network here, and the slice shipped
with the reference is 2,885 all-BENIGN rows — SURVEY.md §0 — on which the binary metrics degenerate).

The generated CSV has the real file's schema quirks: 79 columns, the duplicated ``Fwd Header Length`` header,
headers with leading blanks, ``Infinity`` cells in ``Flow Bytes/s`` / ``Flow Packets/s`` for zero-duration flows, and
both ``BENIGN`` and ``DDoS`` labels with separable (but overlapping) feature distributions, so that the classifier
has something to learn and precision/recall are meaningful.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import pandas as pd

# header of the CICIDS2017 "MachineLearningCVE" files, in file order (note the duplicate and the leading blanks)
CICIDS_HEADER = [
    "Destination Port", "Flow Duration", "Total Fwd Packets", "Total Backward Packets", "Total Length of Fwd Packets",
    "Total Length of Bwd Packets", "Fwd Packet Length Max", "Fwd Packet Length Min", "Fwd Packet Length Mean",
    "Fwd Packet Length Std", "Bwd Packet Length Max", "Bwd Packet Length Min", "Bwd Packet Length Mean",
    "Bwd Packet Length Std", "Flow Bytes/s", "Flow Packets/s", "Flow IAT Mean", "Flow IAT Std", " Flow IAT Max",
    "Flow IAT Min", "Fwd IAT Total", "Fwd IAT Mean", "Fwd IAT Std", "Fwd IAT Max", " Fwd IAT Min", "Bwd IAT Total",
    "Bwd IAT Mean", "Bwd IAT Std", "Bwd IAT Max", " Bwd IAT Min", "Fwd PSH Flags", "Bwd PSH Flags", "Fwd URG Flags",
    "Bwd URG Flags", "Fwd Header Length", "Bwd Header Length", "Fwd Packets/s", " Bwd Packets/s",
    " Min Packet Length", " Max Packet Length", "Packet Length Mean", "Packet Length Std", "Packet Length Variance",
    "FIN Flag Count", " SYN Flag Count", "RST Flag Count", "PSH Flag Count", "ACK Flag Count", "URG Flag Count",
    "CWE Flag Count", "ECE Flag Count", "Down/Up Ratio", "Average Packet Size", "Avg Fwd Segment Size",
    "Avg Bwd Segment Size", "Fwd Header Length", "Fwd Avg Bytes/Bulk", "Fwd Avg Packets/Bulk", "Fwd Avg Bulk Rate",
    "Bwd Avg Bytes/Bulk", "Bwd Avg Packets/Bulk", "Bwd Avg Bulk Rate", "Subflow Fwd Packets", "Subflow Fwd Bytes",
    "Subflow Bwd Packets", "Subflow Bwd Bytes", "Init_Win_bytes_forward", "Init_Win_bytes_backward",
    "act_data_pkt_fwd", "min_seg_size_forward", "Active Mean", "Active Std", "Active Max", "Active Min", "Idle Mean",
    "Idle Std", "Idle Max", "Idle Min", "Label",
]


def synth_flows(n_rows: int, ddos_fraction: float = 0.57, seed: int = 0, inf_rows: int = 6) -> pd.DataFrame:
    """Returns a DataFrame with the ten columns the model reads plus ``Label`` (all other columns are zeros and only
    added when writing a CSV).  Default class balance ≈ the Friday-afternoon DDoS capture (128,027 / 225,745)."""
    rng = np.random.default_rng(seed)
    is_ddos = rng.random(n_rows) < ddos_fraction
    n = n_rows
    well_known = rng.random(n) < 0.7
    benign_port = np.where(well_known, rng.choice([53, 443, 80, 22, 123, 8080, 3389], n), rng.integers(1024, 65535, n))
    port = np.where(is_ddos, 80, benign_port).astype(np.int64)
    dur = np.where(is_ddos, rng.lognormal(14.5, 1.6, n), rng.lognormal(10.5, 3.0, n)).astype(np.int64) + 1
    fwd_p = np.where(is_ddos, rng.integers(2, 9, n), rng.integers(1, 40, n)).astype(np.int64)
    bwd_p = np.where(is_ddos, rng.integers(0, 7, n), rng.integers(0, 45, n)).astype(np.int64)
    fwd_max = np.where(is_ddos, rng.choice([6, 20, 26], n), rng.integers(0, 1500, n)).astype(np.int64)
    fwd_min = np.minimum(fwd_max, np.where(is_ddos, rng.choice([0, 6], n), rng.integers(0, 90, n))).astype(np.int64)
    fwd_len = (fwd_p * (fwd_max + fwd_min) // 2).astype(np.int64)
    bwd_len = np.where(is_ddos, rng.choice([0, 11595, 11601, 11607], n), bwd_p * rng.integers(0, 1460, n)).astype(np.int64)
    secs = dur / 1e6
    bps = (fwd_len + bwd_len) / secs
    pps = (fwd_p + bwd_p) / secs
    df = pd.DataFrame({
        "Destination Port": port, "Flow Duration": dur, "Total Fwd Packets": fwd_p, "Total Backward Packets": bwd_p,
        "Total Length of Fwd Packets": fwd_len, "Total Length of Bwd Packets": bwd_len,
        "Fwd Packet Length Max": fwd_max, "Fwd Packet Length Min": fwd_min,
        "Flow Bytes/s": np.round(bps, 8), "Flow Packets/s": np.round(pps, 8),
        "Label": np.where(is_ddos, "DDoS", "BENIGN"),
    })
    if inf_rows and n_rows > inf_rows:
        idx = rng.choice(n_rows, inf_rows, replace=False)
        df.loc[idx[: inf_rows // 2], "Flow Bytes/s"] = np.inf
        df.loc[idx[inf_rows // 2:], "Flow Packets/s"] = np.inf
    return df


def write_synthetic_csv(path: str, n_rows: int = 22575, ddos_fraction: float = 0.57, seed: int = 0) -> str:
    """Writes a full 79-column CICIDS2017-style CSV (duplicate header and ``Infinity`` spelling included)."""
    core = synth_flows(n_rows, ddos_fraction, seed)
    cols = []
    seen_fwd_header = False
    data = {}
    for i, name in enumerate(CICIDS_HEADER):
        key = f"c{i}"
        cols.append(key)
        if name in core.columns:
            data[key] = core[name]
        elif name == "Fwd Header Length":
            data[key] = core["Total Fwd Packets"] * 20
            seen_fwd_header = True
        else:
            data[key] = np.zeros(n_rows, dtype=np.int64)
    assert seen_fwd_header
    out = pd.DataFrame(data, columns=cols)
    with open(path, "w", newline="") as f:
        f.write(",".join(CICIDS_HEADER) + "\n")
        out.to_csv(f, header=False, index=False, na_rep="NaN")
    # pandas writes inf as "inf"; the real capture spells it "Infinity"
    txt = open(path).read().replace(",inf,", ",Infinity,")
    open(path, "w").write(txt)
    return path


# Measured on 22,575 synthetic flow sentences tokenised with the synthetic vocabulary (tools: data/tokenizer.py):
# 83.1 tokens per sentence (78..88), 2,232 DISTINCT token ids in total, the 100 most frequent carrying 86 % of all
# tokens -- flow records are ten numbers in a fixed sentence frame.  The HF tokenizer of the reference arm gives
# 84.3 (82..86) on the same sentences.
FLOW_DISTINCT_TOKENS = 2232
FLOW_ZIPF_EXPONENT = 1.3      # sum_{k<=100} k^-1.3 / sum_{k<=2232} k^-1.3 = 0.86


def synth_token_batch(batch: int, seq: int, vocab_size: int, seed: int = 0, mean_len: int = 84, len_std: float = 2.0,
                      distinct: int = FLOW_DISTINCT_TOKENS):
    """Token-level synthetic batch shaped like tokenised flow sentences: ~84 real tokens of ``seq`` (std 2), ids drawn
    from a fixed subset of ``distinct`` vocabulary entries with a Zipf(1.3) rank-frequency law (see the measurements
    above), [CLS]=101 first, [SEP]=102 last, [PAD]=0 after.  Returns (input_ids [B,S] int32, lens [B] int32, labels
    [B] int32).  ``distinct <= 0`` draws ids uniformly from the whole vocabulary (the round-1 behaviour: every step
    touches ~1,300 new embedding rows, which no tokenised flow data does)."""
    import torch

    g = torch.Generator().manual_seed(seed)
    lens = torch.clamp((torch.randn(batch, generator=g) * len_std + mean_len).round().long(), 8, seq)
    lo = min(1000, max(vocab_size - 1, 1))
    if distinct and distinct > 0 and vocab_size - lo > 1:
        n = min(distinct, vocab_size - lo)
        used = lo + torch.randperm(vocab_size - lo, generator=torch.Generator().manual_seed(20250805))[:n]
        w = torch.arange(1, n + 1, dtype=torch.float64) ** -FLOW_ZIPF_EXPONENT
        ids = used[torch.multinomial(w, batch * seq, replacement=True, generator=g)].view(batch, seq)
    else:
        ids = torch.randint(lo, max(vocab_size, lo + 1), (batch, seq), generator=g)
    pos = torch.arange(seq)[None, :]
    ids[:, 0] = 101
    ids[pos == (lens[:, None] - 1)] = 102
    ids[pos >= lens[:, None]] = 0
    labels = torch.randint(0, 2, (batch,), generator=g)
    return ids.int(), lens.int(), labels.int()
