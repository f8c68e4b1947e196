"""The notebooks' evaluation cells on the GPU: ``roc_curve``, ``auc``, ``precision_recall_curve``, ``confusion_matrix``.

Reference (Python-Tensorflow-2.0-Keras-Fraud-Detection-Autoencoder.ipynb cells 23, 24, 27)::

    false_pos_rate, true_pos_rate, thresholds = roc_curve(error_df.True_class, error_df.Reconstruction_error)
    roc_auc = auc(false_pos_rate, true_pos_rate)
    precision_rt, recall_rt, threshold_rt = precision_recall_curve(error_df.True_class, error_df.Reconstruction_error)
    conf_matrix = confusion_matrix(error_df.True_class, pred_y)        # pred_y = mse > threshold_fixed

The heavy step — sorting the scores and accumulating true / false positives per distinct threshold — is ``b2_threshold_sweep``
(csrc/metrics.cu); what is left (ratios, the "drop collinear points" filters, the trapezoid rule) is elementwise work on the
thresholds.  Semantics follow the scikit-learn installed next to this package (1.9: ``thresholds[0] = inf`` for ``roc_curve``,
``drop_intermediate`` default True for ``roc_curve`` and False for ``precision_recall_curve``); scores are compared as float32,
which is what the scoring kernels produce.  Binary labels only (non-zero = positive), like the notebooks' ``True_class``.
"""
from __future__ import annotations

import warnings
from typing import Tuple

import numpy as np
import torch

from . import _ffi
from .keras_compat import _stream_ptr, get_ctx


def _dev(*xs) -> torch.device:
    for x in xs:
        if isinstance(x, torch.Tensor) and x.is_cuda:
            return x.device
    return torch.device("cuda", torch.cuda.current_device())


def _labels(y, dev) -> torch.Tensor:
    t = y if isinstance(y, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(y)))
    return (t.to(dev) != 0).to(torch.uint8).contiguous()


def threshold_sweep(y_true, y_score) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """scikit-learn's ``confusion_matrix_at_thresholds`` (formerly ``_binary_clf_curve``): ``(fps, tps, thresholds)`` as device
    tensors (int64, int64, float32), thresholds decreasing, one per distinct score."""
    dev = _dev(y_score, y_true)
    s = y_score if isinstance(y_score, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(y_score, dtype=np.float32)))
    s = s.to(dev, torch.float32).contiguous()
    y = _labels(y_true, dev)
    n = int(s.numel())
    if n == 0 or y.numel() != n or s.dim() != 1:
        raise ValueError("y_true and y_score must be 1-D and of the same, non-zero length")
    lib = _ffi.lib()
    nbytes = int(lib.b2_threshold_sweep_scratch_bytes(n))
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    thr = torch.empty(n, dtype=torch.float32, device=dev)
    tps = torch.empty(n, dtype=torch.int64, device=dev)
    fps = torch.empty(n, dtype=torch.int64, device=dev)
    cnt = torch.zeros(2, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _ffi.check(lib.b2_threshold_sweep(get_ctx(dev.index), _ffi.ptr(s), _ffi.ptr(y), n, _ffi.ptr(thr), _ffi.ptr(tps), _ffi.ptr(fps),
                                          _ffi.ptr(cnt), _ffi.ptr(scratch), nbytes, _stream_ptr()), "b2_threshold_sweep")
    k, nans = (int(v) for v in cnt.cpu().tolist())
    if nans:
        raise ValueError("Input contains NaN.")
    return fps[:k], tps[:k], thr[:k]


def roc_curve(y_true, y_score, drop_intermediate: bool = True):
    """``sklearn.metrics.roc_curve`` → (fpr, tpr, thresholds) as NumPy arrays (float64, float64, float32-valued float64)."""
    fps, tps, thr = threshold_sweep(y_true, y_score)
    if drop_intermediate and fps.numel() > 2:
        # keep the corners: points where the second difference of fps or of tps is non-zero, plus both ends
        d2 = lambda v: v[2:] - 2 * v[1:-1] + v[:-2]  # noqa: E731
        keep = torch.ones(fps.numel(), dtype=torch.bool, device=fps.device)
        keep[1:-1] = (d2(fps) != 0) | (d2(tps) != 0)
        fps, tps, thr = fps[keep], tps[keep], thr[keep]
    fps = np.concatenate([[0.0], fps.cpu().numpy().astype(np.float64)])
    tps = np.concatenate([[0.0], tps.cpu().numpy().astype(np.float64)])
    thresholds = np.concatenate([[np.inf], thr.cpu().numpy().astype(np.float64)])
    if fps[-1] <= 0:
        warnings.warn("No negative samples in y_true, false positive value should be meaningless")
        fpr = np.full(fps.shape, np.nan)
    else:
        fpr = fps / fps[-1]
    if tps[-1] <= 0:
        warnings.warn("No positive samples in y_true, true positive value should be meaningless")
        tpr = np.full(tps.shape, np.nan)
    else:
        tpr = tps / tps[-1]
    return fpr, tpr, thresholds


def precision_recall_curve(y_true, y_score, drop_intermediate: bool = False):
    """``sklearn.metrics.precision_recall_curve`` → (precision, recall, thresholds), recall decreasing, last point (1, 0)."""
    fps, tps, thr = threshold_sweep(y_true, y_score)
    if drop_intermediate and fps.numel() > 2:
        keep = torch.ones(fps.numel(), dtype=torch.bool, device=fps.device)
        keep[1:-1] = (tps[1:-1] != tps[:-2]) | (tps[2:] != tps[1:-1])
        fps, tps, thr = fps[keep], tps[keep], thr[keep]
    fps, tps = fps.cpu().numpy().astype(np.float64), tps.cpu().numpy().astype(np.float64)
    ps = tps + fps
    precision = np.zeros_like(tps)
    np.divide(tps, ps, out=precision, where=ps != 0)
    if tps[-1] == 0:
        warnings.warn("No positive class found in y_true, recall is set to one for all thresholds.")
        recall = np.ones_like(tps)
    else:
        recall = tps / tps[-1]
    return (np.concatenate([precision[::-1], [1.0]]), np.concatenate([recall[::-1], [0.0]]), thr.cpu().numpy()[::-1].copy())


def auc(x, y) -> float:
    """``sklearn.metrics.auc``: trapezoid rule over a monotonic x."""
    x, y = np.asarray(x, dtype=np.float64).ravel(), np.asarray(y, dtype=np.float64).ravel()
    if x.shape[0] < 2:
        raise ValueError("At least 2 points are needed to compute area under curve, but x.shape = %s" % (x.shape,))
    dx = np.diff(x)
    direction = 1
    if np.any(dx < 0):
        if np.all(dx <= 0):
            direction = -1
        else:
            raise ValueError("x is neither increasing nor decreasing : {}.".format(x))
    return float(direction * np.trapezoid(y, x))


def confusion_matrix(y_true, y_pred) -> np.ndarray:
    """``sklearn.metrics.confusion_matrix`` for two classes → ``[[tn, fp], [fn, tp]]`` (int64); e.g. ``y_pred = flag`` from
    ``Model.score`` (mse > threshold_fixed, nb2.0:1129)."""
    dev = _dev(y_true, y_pred)
    yt, yp = _labels(y_true, dev), _labels(y_pred, dev)
    if yt.numel() != yp.numel():
        raise ValueError("y_true and y_pred differ in length")
    cm = torch.zeros(4, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _ffi.check(_ffi.lib().b2_confusion_matrix(get_ctx(dev.index), _ffi.ptr(yt), _ffi.ptr(yp), int(yt.numel()), _ffi.ptr(cm), _stream_ptr()),
                   "b2_confusion_matrix")
    return cm.cpu().numpy().reshape(2, 2)
