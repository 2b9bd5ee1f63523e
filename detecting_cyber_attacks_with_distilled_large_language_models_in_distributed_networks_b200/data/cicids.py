"""CICIDS2017 flow records -> English sentences + binary labels.

Behavioural parity with the reference's L0 data layer (client1.py:68-93): same ten columns in the same order, the
same sentence template (byte-for-byte: integer columns print as ``80``, float columns as Python ``repr`` of the
float64), ``±inf -> NaN -> column mean`` cleaning, ``df.sample(frac, random_state=seed)`` sub-sampling and
``label = 1 iff Label == 'DDoS'``.  The per-client seed is ``41 + client_id`` (client1.py:89 uses 42,
client2.py:84 uses 43).  A disjoint-shard option is an extension.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import pandas as pd

FEATURE_COLUMNS: Tuple[str, ...] = (
    "Destination Port", "Flow Duration", "Total Fwd Packets", "Total Backward Packets",
    "Total Length of Fwd Packets", "Total Length of Bwd Packets", "Fwd Packet Length Max",
    "Fwd Packet Length Min", "Flow Bytes/s", "Flow Packets/s",
)

_TEMPLATE = (
    "Destination port is {0}. "
    "Flow duration is {1} microseconds. "
    "Total forward packets are {2}. "
    "Total backward packets are {3}. "
    "Total length of forward packets is {4} bytes. "
    "Total length of backward packets is {5} bytes. "
    "Maximum forward packet length is {6}. "
    "Minimum forward packet length is {7}. "
    "Flow bytes per second is {8}. "
    "Flow packets per second is {9}."
)

POSITIVE_LABEL = "DDoS"


def client_seed(client_id: int) -> int:
    """Reference seeds: client 1 -> 42, client 2 -> 43 (client1.py:89, client2.py:84)."""
    return 41 + int(client_id)


def features_to_text(row) -> str:
    """One flow -> one sentence.  ``row`` is anything indexable by column name (pandas Series, dict)."""
    return _TEMPLATE.format(*(row[c] for c in FEATURE_COLUMNS))


def _fmt(v) -> str:
    # what an f-string prints for the numpy scalars pandas hands out: ints as ints, floats via repr(float)
    if isinstance(v, (np.integer, int)) and not isinstance(v, bool):
        return str(int(v))
    if isinstance(v, (np.floating, float)):
        return repr(float(v))
    return str(v)


def frame_to_texts(df: pd.DataFrame) -> List[str]:
    """Vectorised equivalent of ``df.apply(features_to_text, axis=1)`` (same strings, ~50x faster)."""
    cols = [df[c].tolist() for c in FEATURE_COLUMNS]  # tolist() yields python ints / floats per column dtype
    strs = [[_fmt(v) for v in col] for col in cols]
    return [_TEMPLATE.format(*vals) for vals in zip(*strs)]


def clean_frame(df: pd.DataFrame) -> pd.DataFrame:
    """``replace(±inf, NaN)`` then ``fillna(mean of numeric columns)`` (client1.py:87-88)."""
    df = df.replace([np.inf, -np.inf], np.nan)
    return df.fillna(df.mean(numeric_only=True))


def labels_from_frame(df: pd.DataFrame) -> List[int]:
    return [1 if x == POSITIVE_LABEL else 0 for x in df["Label"].tolist()]


def preprocess_data(file_path: str, data_fraction: float = 0.1, seed: int = 42,
                    shard: Optional[Tuple[int, int]] = None) -> Tuple[List[str], List[int]]:
    """Reference ``preprocess_data`` (client1.py:84-93).  ``shard=(index, count)`` (extension) first restricts the
    file to a disjoint ``index``-th of ``count`` interleaved shards so clients do not overlap."""
    df = pd.read_csv(file_path)
    df = clean_frame(df)
    if shard is not None:
        idx, cnt = shard
        df = df.iloc[idx::cnt]
    df = df.sample(frac=data_fraction, random_state=seed)
    return frame_to_texts(df), labels_from_frame(df)
