"""report/metrics.py against the library calls the reference makes (client1.py:143-146, 167-193, 339-350):
sklearn ``precision_recall_fscore_support(average='binary')``, ``confusion_matrix``, ``roc_curve`` / ``auc`` /
``precision_recall_curve``, and the one-row metrics CSV."""
import warnings

import numpy as np
import pytest

from detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200.report.metrics import (
    CSV_COLUMNS, EvalResult, binary_prf_from_confusion, pr_curve_points, roc_curve_points, save_metrics)

skm = pytest.importorskip("sklearn.metrics")


def _cases():
    rng = np.random.default_rng(0)
    yield rng.integers(0, 2, 500), rng.integers(0, 2, 500)          # generic
    yield np.zeros(64, dtype=np.int64), np.zeros(64, dtype=np.int64)  # the shipped CSV: every flow BENIGN, none predicted
    yield rng.integers(0, 2, 100), np.zeros(100, dtype=np.int64)     # nothing predicted positive: precision 0/0
    yield np.zeros(100, dtype=np.int64), rng.integers(0, 2, 100)     # no positive label: recall 0/0
    yield np.ones(40, dtype=np.int64), np.ones(40, dtype=np.int64)   # all positive, all right


@pytest.mark.parametrize("case", list(_cases()), ids=["generic", "all_benign", "no_pos_pred", "no_pos_label", "all_pos"])
def test_confusion_derived_metrics_equal_sklearn(case):
    y, pred = case
    cm = np.zeros((2, 2), dtype=np.int64)
    np.add.at(cm, (y, pred), 1)                                       # what the eval kernel accumulates on the GPU
    assert np.array_equal(cm, skm.confusion_matrix(y, pred, labels=[0, 1]))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")                               # sklearn warns on 0/0 and returns 0.0
        p, r, f, _ = skm.precision_recall_fscore_support(y, pred, average="binary")
    assert binary_prf_from_confusion(cm) == pytest.approx((p, r, f), abs=1e-12)
    res = EvalResult.from_confusion(cm, loss_sum=3.0, n_batches=4)
    assert res.accuracy == pytest.approx(100.0 * float((y == pred).mean()))   # percent, client1.py:143
    assert res.loss == 0.75                                                    # mean of batch means, client1.py:144
    assert res[0] == res.accuracy and res[4] == res.f1 and len(res.as_tuple()) == 8


def test_roc_and_pr_curves_equal_sklearn():
    rng = np.random.default_rng(1)
    y = rng.integers(0, 2, 300)
    s = np.round(rng.random(300) * 0.6 + 0.3 * y, 2)                  # ties on purpose
    fpr, tpr, auc = roc_curve_points(y, s)
    f2, t2, _ = skm.roc_curve(y, s, drop_intermediate=False)
    assert np.allclose(fpr, f2) and np.allclose(tpr, t2)
    assert auc == pytest.approx(skm.auc(f2, t2), abs=1e-12)
    prec, rec = pr_curve_points(y, s)
    p2, r2, _ = skm.precision_recall_curve(y, s)
    assert np.allclose(prec, p2) and np.allclose(rec, r2)


def test_metrics_csv_is_the_reference_format(tmp_path):
    res = EvalResult.from_confusion([[50, 3], [7, 40]], loss_sum=1.2345678901234, n_batches=7)
    path = tmp_path / "client1_local_metrics.csv"
    save_metrics(res, str(path))
    lines = path.read_text().splitlines()
    assert lines[0] == ",".join(CSV_COLUMNS) == "Accuracy,Loss,Precision,Recall,F1-Score"
    assert len(lines) == 2
    vals = [float(x) for x in lines[1].split(",")]
    assert vals == pytest.approx([res.accuracy, res.loss, res.precision, res.recall, res.f1], rel=1e-15)
