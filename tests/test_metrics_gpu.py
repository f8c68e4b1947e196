"""K7 parity: the notebooks' evaluation metrics (roc_curve, auc, precision_recall_curve, confusion_matrix) computed on the GPU
against scikit-learn itself — the library the reference calls (nb2.0 cells 23, 24, 27)."""
import numpy as np
import pytest

import b200_sensor as bs
from oracle import cpu_oracle as orc

pytestmark = pytest.mark.gpu


def _cases():
    g = np.random.default_rng(5)
    for n, pos, kind in [(1, 1.0, "cont"), (2, 0.5, "cont"), (7, 0.4, "ties"), (2047, 0.1, "cont"), (2048, 0.02, "ties"), (2049, 0.5, "cont"),
                         (5000, 0.002, "cont"), (100_000, 0.0017, "mse"), (300_001, 0.3, "ties"), (1 << 20, 0.01, "mse")]:
        y = (g.random(n) < pos).astype(np.int64)
        if kind == "cont":
            s = g.normal(size=n)
        elif kind == "ties":
            s = g.integers(-5, 6, n) * 0.5           # heavy ties, negative and zero scores, -0.0 among them
            s[::13] = -0.0
        else:
            s = np.abs(g.normal(size=n)) ** 2 + y * g.uniform(0, 8, n)   # reconstruction-error-like: positives score higher
        yield n, y, s.astype(np.float32)


def test_threshold_sweep_roc_pr_auc_match_sklearn(cuda_device):
    for n, y, s in _cases():
        if y.min() == y.max():
            y[0] = 1 - y[0] if n > 1 else y[0]
        for drop in (True, False):
            fpr, tpr, thr = bs.metrics.roc_curve(y, s, drop_intermediate=drop)
            if n > 1 and y.min() != y.max():
                rf, rt, rthr = orc.sk_roc_curve(y, s, drop_intermediate=drop)
                np.testing.assert_array_equal(fpr, rf)
                np.testing.assert_array_equal(tpr, rt)
                np.testing.assert_array_equal(thr, rthr.astype(np.float64))
                assert bs.metrics.auc(fpr, tpr) == orc.sk_auc(rf, rt)
            p, r, t = bs.metrics.precision_recall_curve(y, s, drop_intermediate=drop)
            if y.max() == 1:
                rp, rr, rt2 = orc.sk_precision_recall_curve(y, s, drop_intermediate=drop)
                np.testing.assert_array_equal(p, rp)
                np.testing.assert_array_equal(r, rr)
                np.testing.assert_array_equal(t, rt2)


def test_confusion_matrix_and_model_flags(cuda_device):
    g = np.random.default_rng(9)
    for n in (1, 31, 4097, 1_000_003):
        yt, yp = g.integers(0, 2, n), g.integers(0, 2, n)
        np.testing.assert_array_equal(bs.metrics.confusion_matrix(yt, yp), orc.sk_confusion_matrix(yt, yp))
    assert bs.metrics.confusion_matrix(np.zeros(0), np.zeros(0)).tolist() == [[0, 0], [0, 0]]


def test_nan_scores_are_refused(cuda_device):
    with pytest.raises(ValueError):
        bs.metrics.roc_curve(np.array([0, 1, 1]), np.array([0.1, np.nan, 0.3], np.float32))
