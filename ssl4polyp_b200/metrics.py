"""Evaluation metrics of the fine-tune loop, restated without scikit-learn (numpy on the host: B x n_class scalars per
epoch, not a device path).

Reference: classification/metrics/performance.py:100-231 (meanF1Score, meanPrecision, meanRecall, meanAUROC,
meanBalancedAccuracy, meanAUPRC) and metrics/thresholds.py:68-110 (compute_youden_j_threshold).  The reference calls
sklearn's roc_auc_score / average_precision_score / balanced_accuracy_score / roc_curve; tests/test_host.py checks these
restatements against sklearn on seeded inputs with ties.
"""
from __future__ import annotations

import math

import numpy as np

_SMOOTH = 1e-8          # performance.py:101 (smooth=1e-8 in every mean* module)


def _np(a):
    if hasattr(a, "detach"):
        a = a.detach().float().cpu().numpy()
    return np.asarray(a)


def positive_scores(probs) -> np.ndarray:
    """[N] positive-class score from [N] or [N,2] probabilities (performance.py:39-58)."""
    p = _np(probs).astype(np.float64)
    return p if p.ndim == 1 else p[:, -1]


def labels_from(probs, tau=None) -> np.ndarray:
    """performance.py:73-97: binary with tau -> score >= tau, otherwise argmax."""
    p = _np(probs)
    if tau is not None and (p.ndim == 1 or p.shape[1] == 2):
        return (positive_scores(p) >= float(tau)).astype(np.int64)
    if p.ndim == 1:
        return (p >= 0.5).astype(np.int64)
    return p.argmax(axis=1).astype(np.int64)


def _roc_points(y: np.ndarray, s: np.ndarray):
    """Cumulative (fp, tp, thresholds) at each distinct score, descending - sklearn's _binary_clf_curve."""
    order = np.argsort(-s, kind="mergesort")
    s, y = s[order], y[order]
    cut = np.r_[np.nonzero(np.diff(s))[0], y.size - 1]
    tp = np.cumsum(y)[cut].astype(np.float64)
    fp = (1 + cut - tp).astype(np.float64)
    return fp, tp, s[cut]


def binary_auroc(y, s) -> float:
    y, s = _np(y).astype(np.int64), _np(s).astype(np.float64)
    if np.unique(y).size < 2:
        return float("nan")          # performance.py:180-186 (warns, returns NaN)
    fp, tp, _ = _roc_points(y, s)
    fpr = np.r_[0.0, fp / fp[-1]]
    tpr = np.r_[0.0, tp / tp[-1]]
    return float(np.trapezoid(tpr, fpr))


def binary_auprc(y, s) -> float:
    """Average precision: sum over thresholds of (R_n - R_{n-1}) * P_n."""
    y, s = _np(y).astype(np.int64), _np(s).astype(np.float64)
    if y.sum() == 0:
        return 0.0
    fp, tp, _ = _roc_points(y, s)
    precision = tp / (tp + fp)
    recall = tp / tp[-1]
    return float(np.sum(np.diff(np.r_[0.0, recall]) * precision))


def auroc(probs, targets, n_class: int) -> float:
    """performance.py:158-195: binary uses column 1; otherwise one-vs-rest macro average."""
    p, y = _np(probs).astype(np.float64), _np(targets).astype(np.int64)
    if np.unique(y).size < 2:
        return float("nan")
    if n_class == 2:
        return binary_auroc(y, positive_scores(p))
    return float(np.mean([binary_auroc((y == c).astype(np.int64), p[:, c]) for c in range(n_class)]))


def auprc(probs, targets, n_class: int) -> float:
    p, y = _np(probs).astype(np.float64), _np(targets).astype(np.int64)
    if n_class == 2:
        return binary_auprc(y, positive_scores(p))
    return float(np.mean([binary_auprc((y == c).astype(np.int64), p[:, c]) for c in range(n_class)]))


def _per_class(probs, targets, n_class, tau, denom):
    lab, y = labels_from(probs, tau), _np(targets).astype(np.int64)
    total = 0.0
    for c in range(n_class):
        m1, m2 = lab == c, y == c
        inter = float(np.sum(m1 & m2))
        total += denom(inter, float(m1.sum()), float(m2.sum()))
    return total / n_class


def mean_f1(probs, targets, n_class: int, tau=None) -> float:
    return _per_class(probs, targets, n_class, tau, lambda i, a, b: 2.0 * (i + _SMOOTH) / (a + b + _SMOOTH))


def mean_precision(probs, targets, n_class: int, tau=None) -> float:
    return _per_class(probs, targets, n_class, tau, lambda i, a, b: (i + _SMOOTH) / (a + _SMOOTH))


def mean_recall(probs, targets, n_class: int, tau=None) -> float:
    return _per_class(probs, targets, n_class, tau, lambda i, a, b: (i + _SMOOTH) / (b + _SMOOTH))


def balanced_accuracy(probs, targets, n_class: int, tau=None) -> float:
    """Mean recall over the classes PRESENT in targets (sklearn's definition, performance.py:198-210)."""
    lab, y = labels_from(probs, tau), _np(targets).astype(np.int64)
    rec = [float(np.mean(lab[y == c] == c)) for c in np.unique(y)]
    return float(np.mean(rec)) if rec else float("nan")


def mcc(probs, targets, tau=None) -> float:
    lab, y = labels_from(probs, tau), _np(targets).astype(np.int64)
    tp = float(np.sum((lab == 1) & (y == 1))); tn = float(np.sum((lab == 0) & (y == 0)))
    fp = float(np.sum((lab == 1) & (y == 0))); fn = float(np.sum((lab == 0) & (y == 1)))
    den = math.sqrt((tp + fp) * (tp + fn) * (tn + fp) * (tn + fn))
    return (tp * tn - fp * fn) / den if den > 0 else 0.0


def youden_j_threshold(probs, targets) -> float:
    """thresholds.py:68-110: tau on the positive-class probability maximising tpr - fpr (first maximum, thresholds
    descending, the leading (0,0) point carrying +inf -> nextafter(max score, 1))."""
    y, s = _np(targets).astype(np.int64), positive_scores(probs)
    if s.size == 0:
        raise ValueError("Cannot compute threshold on empty logits tensor")
    if np.unique(y).size < 2:
        raise ValueError("Youden's J threshold requires both positive and negative samples")
    fp, tp, thr = _roc_points(y, s)
    # sklearn.roc_curve drops collinear interior points (drop_intermediate=True) before the argmax
    fps, tps, thr = np.r_[0.0, fp], np.r_[0.0, tp], np.r_[np.inf, thr]
    if fps.size > 2:
        keep = np.r_[True, np.logical_or(np.diff(fps, 2), np.diff(tps, 2)), True]
        fps, tps, thr = fps[keep], tps[keep], thr[keep]
    j = tps / tps[-1] - fps / fps[-1]
    tau = float(thr[int(np.argmax(j))])
    if math.isinf(tau):
        tau = float(np.nextafter(s.max(), 1.0))
    return tau
