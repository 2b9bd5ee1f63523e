"""Evaluation metrics with the reference's exact definitions, and the metrics CSV writer.

Reference ``evaluate_model`` (client1.py:118-150): accuracy in PERCENT (``100*correct/total``), loss = mean of the
per-batch mean losses (``total_loss/len(loader)``, not a per-sample mean — SURVEY.md Q8), and sklearn
``precision_recall_fscore_support(average='binary')`` / ``confusion_matrix``.  Here the 2x2 confusion matrix is
accumulated on the GPU and precision/recall/F1 are derived from it with sklearn's zero-division convention (0.0).
``save_metrics`` writes the same one-row CSV (header ``Accuracy,Loss,Precision,Recall,F1-Score``, client1.py:339-350).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

CSV_COLUMNS = ("Accuracy", "Loss", "Precision", "Recall", "F1-Score")


def binary_prf_from_confusion(cm) -> Tuple[float, float, float]:
    """cm[true, pred]; positive class = 1.  Zero division -> 0.0 (sklearn's default behaviour)."""
    cm = np.asarray(cm, dtype=np.int64)
    tp, fp, fn = int(cm[1, 1]), int(cm[0, 1]), int(cm[1, 0])
    precision = tp / (tp + fp) if (tp + fp) else 0.0
    recall = tp / (tp + fn) if (tp + fn) else 0.0
    f1 = 2 * precision * recall / (precision + recall) if (precision + recall) else 0.0
    return float(precision), float(recall), float(f1)


@dataclass
class EvalResult:
    accuracy: float            # percent
    loss: float                # mean of batch-mean losses
    precision: float
    recall: float
    f1: float
    confusion: np.ndarray      # [2,2] int64, [true, pred]
    labels: List[int] = field(default_factory=list)
    probs: List[float] = field(default_factory=list)

    def as_tuple(self):
        """The reference's 8-tuple (client1.py:150)."""
        return (self.accuracy, self.loss, self.precision, self.recall, self.f1, self.confusion, self.labels,
                self.probs)

    def __getitem__(self, i):
        return self.as_tuple()[i]

    @classmethod
    def from_confusion(cls, cm, loss_sum: float, n_batches: int, labels=None, probs=None) -> "EvalResult":
        cm = np.asarray(cm, dtype=np.int64).reshape(2, 2)
        total = int(cm.sum())
        correct = int(cm[0, 0] + cm[1, 1])
        p, r, f = binary_prf_from_confusion(cm)
        return cls(100.0 * correct / total if total else 0.0, loss_sum / n_batches if n_batches else 0.0, p, r, f,
                   cm, list(labels) if labels is not None else [], list(probs) if probs is not None else [])


def save_metrics(metrics, filename: str) -> None:
    """One-row CSV exactly like the reference (pandas ``to_csv(index=False)`` float formatting)."""
    import pandas as pd

    row = {c: metrics[i] for i, c in enumerate(CSV_COLUMNS)}
    pd.DataFrame([row]).to_csv(filename, index=False)


def roc_curve_points(labels: Sequence[int], probs: Sequence[float]):
    """ROC curve + AUC (the reference defines but never calls ``plot_roc_curve``, client1.py:167-180; offered here as
    an optional output).  Plain numpy, same thresholds convention as sklearn (descending distinct scores)."""
    y = np.asarray(labels, dtype=np.int64)
    s = np.asarray(probs, dtype=np.float64)
    order = np.argsort(-s, kind="mergesort")
    y, s = y[order], s[order]
    distinct = np.where(np.diff(s))[0]
    idx = np.r_[distinct, y.size - 1]
    tps = np.cumsum(y)[idx]
    fps = 1 + idx - tps
    tps, fps = np.r_[0, tps], np.r_[0, fps]
    tpr = tps / tps[-1] if tps[-1] else np.zeros_like(tps, dtype=np.float64)
    fpr = fps / fps[-1] if fps[-1] else np.zeros_like(fps, dtype=np.float64)
    auc = float(np.trapezoid(tpr, fpr)) if hasattr(np, "trapezoid") else float(np.trapz(tpr, fpr))
    return fpr, tpr, auc


def pr_curve_points(labels: Sequence[int], probs: Sequence[float]):
    y = np.asarray(labels, dtype=np.int64)
    s = np.asarray(probs, dtype=np.float64)
    order = np.argsort(-s, kind="mergesort")
    y, s = y[order], s[order]
    distinct = np.where(np.diff(s))[0]
    idx = np.r_[distinct, y.size - 1]
    tps = np.cumsum(y)[idx]
    fps = 1 + idx - tps
    precision = tps / np.maximum(tps + fps, 1)
    recall = tps / tps[-1] if tps[-1] else np.ones_like(tps, dtype=np.float64)
    return np.r_[precision[::-1], 1.0], np.r_[recall[::-1], 0.0]
