"""60/20/20 train/val/test split with the reference's exact index selection.

Reference: two chained ``sklearn.model_selection.train_test_split`` calls with the same seed
(client1.py:365-366): ``test_size=0.4`` then ``test_size=0.5`` of the remainder.  sklearn is a cold-path
dependency here (the split happens once per client run), so it is simply called; ``split_indices`` returns index
arrays so the tokenised tensors can be sliced without copying Python lists around.
"""
from __future__ import annotations

from typing import Sequence, Tuple

import numpy as np


def split_indices(n: int, seed: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    from sklearn.model_selection import train_test_split

    idx = np.arange(n)
    train, temp = train_test_split(idx, test_size=0.4, random_state=seed)
    val, test = train_test_split(temp, test_size=0.5, random_state=seed)
    return train, val, test


def split_lists(texts: Sequence, labels: Sequence, seed: int):
    """Same return structure as the reference's six lists (X_train, X_val, X_test, y_train, y_val, y_test)."""
    tr, va, te = split_indices(len(texts), seed)
    pick = lambda seq, ix: [seq[i] for i in ix]  # noqa: E731
    return (pick(texts, tr), pick(texts, va), pick(texts, te), pick(labels, tr), pick(labels, va), pick(labels, te))
