"""CPU oracle for the metric-state hot path (numpy restatement of the reference algorithms).

TEST INFRASTRUCTURE ONLY. Nothing in the product package (`torchmetrics_b200/`) imports this
module; only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference`
legs of `bench.py` may call it, and there only as the checker or as the timed CPU baseline.

Every function restates one reference function (path:line relative to /root/reference/src/torchmetrics)
in plain numpy. Integer outputs must be bit-identical to the reference; the float32 steps
(`_safe_divide`, `trapz`, `tps` cumsum in fp32) follow the reference's dtypes.

Parity pinning: `tests/test_oracle_golden.py` checks every function here against fixtures in
`tests/golden/*.npz`, which were produced by running the *unmodified* reference in the build
container (`tests/golden/make_golden.py`, committed). So: parity pinned.
"""
from __future__ import annotations

import numpy as np

I64 = np.int64


# --------------------------------------------------------------------------------------
# A1  utilities/data.py:179-207   _bincount
# --------------------------------------------------------------------------------------
def bincount(x: np.ndarray, minlength: int) -> np.ndarray:
    """Occurrences of each value of non-negative integer `x`; like torch.bincount the result grows when
    max(x) >= minlength and a negative entry is an error (utilities/data.py:207)."""
    x = np.asarray(x).astype(I64).ravel()
    if x.size and x.min() < 0:
        raise RuntimeError("bincount only supports 1-d non-negative integral inputs.")
    return np.bincount(x, minlength=minlength).astype(I64)


def _sigmoid32(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.float32)
    one = np.float32(1.0)
    with np.errstate(over="ignore"):
        return (one / (one + np.exp(-x, dtype=np.float32))).astype(np.float32)


def _in_unit_range(p: np.ndarray) -> bool:
    # torch.all((preds >= 0) * (preds <= 1)); NaN fails both comparisons
    return bool(np.all((p >= 0) & (p <= 1)))


# --------------------------------------------------------------------------------------
# A2/A3  functional/classification/confusion_matrix.py:306-337
# --------------------------------------------------------------------------------------
def multiclass_confusion_matrix_format(preds, target, ignore_index=None):
    """argmax over dim 1 when preds has one more dim than target, flatten, drop ignored targets
    (confusion_matrix.py:306-330)."""
    preds = np.asarray(preds)
    target = np.asarray(target)
    if preds.ndim == target.ndim + 1:
        preds = preds.argmax(axis=1)
    preds = preds.reshape(-1)
    target = target.reshape(-1)
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    return preds, target


def multiclass_confusion_matrix_update(preds, target, num_classes: int) -> np.ndarray:
    """key = target*C + pred, histogram over C*C keys, rows = target, cols = pred
    (confusion_matrix.py:333-337). An out-of-range label is an error, as the reference's
    reshape(C, C) of an over-long bincount is."""
    key = np.asarray(target).astype(I64) * num_classes + np.asarray(preds).astype(I64)
    bins = bincount(key, num_classes * num_classes)
    if bins.size != num_classes * num_classes:
        raise RuntimeError(f"shape '[{num_classes}, {num_classes}]' is invalid for input of size {bins.size}")
    return bins.reshape(num_classes, num_classes)


# --------------------------------------------------------------------------------------
# A4  confusion_matrix.py:118-153 (binary), :486-526 (multilabel)
# --------------------------------------------------------------------------------------
def binary_confusion_matrix_format(preds, target, threshold=0.5, ignore_index=None):
    preds = np.asarray(preds).reshape(-1)
    target = np.asarray(target).reshape(-1)
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    if np.issubdtype(preds.dtype, np.floating):
        if not _in_unit_range(preds):
            preds = _sigmoid32(preds)
        preds = preds > np.float32(threshold)  # strict, confusion_matrix.py:144
    return preds, target


def binary_confusion_matrix_update(preds, target) -> np.ndarray:
    key = (np.asarray(target).astype(I64) * 2 + np.asarray(preds).astype(I64))
    return bincount(key, 4).reshape(2, 2)


def multilabel_confusion_matrix_format(preds, target, num_labels, threshold=0.5, ignore_index=None):
    preds = np.asarray(preds)
    target = np.asarray(target)
    if np.issubdtype(preds.dtype, np.floating):
        if not _in_unit_range(preds):
            preds = _sigmoid32(preds)
        preds = preds > np.float32(threshold)
    preds = np.moveaxis(preds, 1, -1).reshape(-1, num_labels).astype(I64)
    target = np.moveaxis(target, 1, -1).reshape(-1, num_labels).astype(I64)
    if ignore_index is not None:
        ign = target == ignore_index
        preds = preds.copy()
        target = target.copy()
        preds[ign] = -4 * num_labels
        target[ign] = -4 * num_labels
    return preds, target


def multilabel_confusion_matrix_update(preds, target, num_labels) -> np.ndarray:
    key = (2 * target + preds) + 4 * np.arange(num_labels, dtype=I64)
    key = key.reshape(-1)
    key = key[key >= 0]
    return bincount(key, 4 * num_labels).reshape(num_labels, 2, 2)


# --------------------------------------------------------------------------------------
# A6  functional/classification/stat_scores.py:90-131 (binary), :647-681 (multilabel)
# --------------------------------------------------------------------------------------
def binary_stat_scores_format(preds, target, threshold=0.5, ignore_index=None):
    preds = np.asarray(preds)
    target = np.asarray(target)
    if np.issubdtype(preds.dtype, np.floating):
        if not _in_unit_range(preds):
            preds = _sigmoid32(preds)
        preds = preds > np.float32(threshold)  # strict, stat_scores.py:107
    preds = preds.reshape(preds.shape[0], -1)
    target = target.reshape(target.shape[0], -1).astype(I64)
    if ignore_index is not None:
        target = target.copy()
        target[target == ignore_index] = -1
    return preds.astype(I64), target


def _four_counts(preds, target, axes):
    eq = target == preds
    tp = (eq & (target == 1)).sum(axis=axes, dtype=I64)
    fn = (~eq & (target == 1)).sum(axis=axes, dtype=I64)
    fp = (~eq & (target == 0)).sum(axis=axes, dtype=I64)
    tn = (eq & (target == 0)).sum(axis=axes, dtype=I64)
    return tp, fp, tn, fn


def binary_stat_scores_update(preds, target, multidim_average="global"):
    """tp/fp/tn/fn over dims [0,1] (global) or [1] (samplewise) (stat_scores.py:120-131)."""
    return _four_counts(preds, target, (0, 1) if multidim_average == "global" else (1,))


def multilabel_stat_scores_format(preds, target, num_labels, threshold=0.5, ignore_index=None):
    preds = np.asarray(preds)
    target = np.asarray(target)
    if np.issubdtype(preds.dtype, np.floating):
        if not _in_unit_range(preds):
            preds = _sigmoid32(preds)
        preds = preds > np.float32(threshold)  # strict, stat_scores.py:660
    preds = preds.reshape(preds.shape[0], preds.shape[1], -1).astype(I64)
    target = target.reshape(target.shape[0], target.shape[1], -1).astype(I64)
    if ignore_index is not None:
        target = target.copy()
        target[target == ignore_index] = -1
    return preds, target


def multilabel_stat_scores_update(preds, target, multidim_average="global"):
    """Per-label tp/fp/tn/fn over dims [0,-1] (global) or [-1] (stat_scores.py:672-681)."""
    return _four_counts(preds, target, (0, 2) if multidim_average == "global" else (2,))


# --------------------------------------------------------------------------------------
# A5  stat_scores.py:325-419   multiclass (micro and confusion-matrix branches, top_k == 1, global)
# --------------------------------------------------------------------------------------
def multiclass_stat_scores_format(preds, target):
    preds = np.asarray(preds)
    target = np.asarray(target)
    if preds.ndim == target.ndim + 1:
        preds = preds.argmax(axis=1)
    return preds.reshape(preds.shape[0], -1), target.reshape(target.shape[0], -1)


def multiclass_stat_scores_update(preds, target, num_classes, average="macro", ignore_index=None):
    preds = np.asarray(preds).reshape(-1).astype(I64)
    target = np.asarray(target).reshape(-1).astype(I64)
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    if average == "micro":  # stat_scores.py:394-404
        tp = I64((preds == target).sum())
        fp = I64((preds != target).sum())
        fn = fp
        tn = I64(num_classes * preds.size - (fp + fn + tp))
        return tp, fp, tn, fn
    cm = multiclass_confusion_matrix_update(preds, target, num_classes)  # stat_scores.py:405-418
    tp = np.diag(cm).copy()
    fp = cm.sum(0) - tp
    fn = cm.sum(1) - tp
    tn = cm.sum() - (fp + fn + tp)
    return tp, fp, tn, fn


def multiclass_stat_scores_update_samplewise(preds, target, num_classes, top_k=1, multidim_average="global",
                                             ignore_index=None):
    """One-hot branch (stat_scores.py:363-393): used for samplewise averaging or top_k != 1.
    preds: (N, X) labels for top_k == 1, (N, C, X) scores otherwise; target (N, X)."""
    preds = np.asarray(preds).copy()
    target = np.asarray(target).astype(I64).copy()
    ignore_in = (0 <= ignore_index <= num_classes - 1) if ignore_index is not None else None
    if ignore_index is not None and not ignore_in:
        idx = target == ignore_index
        target[idx] = num_classes
        if preds.ndim > target.ndim:
            preds[np.repeat(idx[:, None, :], num_classes, axis=1)] = num_classes
        else:
            preds[idx] = num_classes
    depth = num_classes + 1 if (ignore_index is not None and not ignore_in) else num_classes
    if top_k > 1:
        # select_topk (utilities/data.py:125-149): ones at the k largest scores along dim 1
        order = np.argsort(-preds, axis=1, kind="stable")[:, :top_k]
        oh = np.zeros(preds.shape, dtype=I64)
        np.put_along_axis(oh, order, 1, axis=1)
        preds_oh = np.moveaxis(oh, 1, -1)
    else:
        preds_oh = np.eye(depth, dtype=I64)[preds.astype(I64)]
    target_oh = np.eye(depth, dtype=I64)[target]
    if ignore_index is not None:
        if ignore_in:
            target_oh[target == ignore_index, :] = -1
        else:
            if top_k == 1:
                preds_oh = preds_oh[..., :-1]
            target_oh = target_oh[..., :-1]
            target_oh[target == num_classes, :] = -1
    axes = (0, 1) if multidim_average == "global" else (1,)
    return _four_counts(preds_oh, target_oh, axes)


# --------------------------------------------------------------------------------------
# A7  functional/classification/precision_recall_curve.py:83-91, :162-250  binned binary curve
# --------------------------------------------------------------------------------------
def binary_precision_recall_curve_format(preds, target, ignore_index=None):
    """flatten, drop ignored, sigmoid when any score is outside [0,1] (precision_recall_curve.py:162-187)."""
    preds = np.asarray(preds).reshape(-1)
    target = np.asarray(target).reshape(-1)
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    if not _in_unit_range(preds):
        preds = _sigmoid32(preds)
    return preds, target


def binary_precision_recall_curve_update_loop(preds, target, thresholds) -> np.ndarray:
    """One threshold at a time; positive means target == 1, everything else is a negative
    (precision_recall_curve.py:228-250)."""
    preds = np.asarray(preds)
    pos = np.asarray(target) == 1
    thr = np.asarray(thresholds)
    T = len(thr)
    cm = np.empty((T, 2, 2), dtype=I64)
    for i in range(T):
        above = preds >= thr[i]
        cm[i, 1, 1] = (pos & above).sum()
        cm[i, 0, 1] = (~pos & above).sum()
        cm[i, 1, 0] = (pos & ~above).sum()
    cm[:, 0, 0] = preds.size - cm[:, 0, 1] - cm[:, 1, 0] - cm[:, 1, 1]
    return cm


def binary_precision_recall_curve_update_hist(preds, target, thresholds) -> np.ndarray:
    """Independent second restatement: bin = #{thr <= p} via searchsorted on the SORTED thresholds,
    histogram, suffix-sum. NaN compares false everywhere so it is forced into bin 0.
    Equals the loop variant for target in {0,1} (+ anything else counted as negative)."""
    preds = np.asarray(preds)
    thr = np.asarray(thresholds)
    order = np.argsort(thr, kind="stable")
    sthr = thr[order]
    b = np.searchsorted(sthr, preds, side="right")
    b[np.isnan(preds)] = 0
    pos = np.asarray(target) == 1
    T = len(thr)
    hp = np.bincount(b[pos], minlength=T + 1).astype(I64)
    hn = np.bincount(b[~pos], minlength=T + 1).astype(I64)
    # count with p >= sthr[i]  == samples whose bin index is > i
    tp_s = hp[::-1].cumsum()[::-1][1:]
    fp_s = hn[::-1].cumsum()[::-1][1:]
    tp = np.empty(T, dtype=I64)
    fp = np.empty(T, dtype=I64)
    tp[order] = tp_s
    fp[order] = fp_s
    cm = np.empty((T, 2, 2), dtype=I64)
    cm[:, 1, 1] = tp
    cm[:, 0, 1] = fp
    cm[:, 1, 0] = hp.sum() - tp
    cm[:, 0, 0] = hn.sum() - fp
    return cm


def binary_precision_recall_curve_update_vectorized(preds, target, thresholds) -> np.ndarray:
    """N x T key tensor + bincount (precision_recall_curve.py:210-225); small N only."""
    preds = np.asarray(preds)
    thr = np.asarray(thresholds)
    T = len(thr)
    above = (preds[:, None] >= thr[None, :]).astype(I64)
    key = above + 2 * np.asarray(target).astype(I64)[:, None] + 4 * np.arange(T, dtype=I64)
    return bincount(key.reshape(-1), 4 * T).reshape(T, 2, 2)


# --------------------------------------------------------------------------------------
# A8  precision_recall_curve.py:423-527 (multiclass), :739-793 (multilabel)
# --------------------------------------------------------------------------------------
def _softmax32(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.float32)
    m = x.max(axis=1, keepdims=True)
    e = np.exp(x - m, dtype=np.float32)
    return (e / e.sum(axis=1, keepdims=True, dtype=np.float32)).astype(np.float32)


def multiclass_precision_recall_curve_format(preds, target, num_classes, ignore_index=None, average=None):
    preds = np.asarray(preds)
    target = np.asarray(target).reshape(-1)
    preds = np.swapaxes(preds, 0, 1).reshape(num_classes, -1).T
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    if not _in_unit_range(preds):
        preds = _softmax32(preds)
    if average == "micro":
        preds = preds.reshape(-1)
        target = np.eye(num_classes, dtype=I64)[target.astype(I64)].reshape(-1)
    return preds, target


def multiclass_precision_recall_curve_update_loop(preds, target, num_classes, thresholds) -> np.ndarray:
    preds = np.asarray(preds)
    thr = np.asarray(thresholds)
    T = len(thr)
    toh = np.eye(num_classes, dtype=bool)[np.asarray(target).astype(I64)]
    cm = np.empty((T, num_classes, 2, 2), dtype=I64)
    for i in range(T):
        above = preds >= thr[i]
        cm[i, :, 1, 1] = (toh & above).sum(0)
        cm[i, :, 0, 1] = (~toh & above).sum(0)
        cm[i, :, 1, 0] = (toh & ~above).sum(0)
    cm[:, :, 0, 0] = preds.shape[0] - cm[:, :, 0, 1] - cm[:, :, 1, 0] - cm[:, :, 1, 1]
    return cm


def multilabel_precision_recall_curve_format(preds, target, num_labels, ignore_index=None):
    """Returns preds (M, L) after optional sigmoid, target (M, L) and a boolean `ignored` mask
    (the reference encodes ignored entries as large negative keys, precision_recall_curve.py:760-766)."""
    preds = np.asarray(preds)
    target = np.asarray(target)
    preds = np.swapaxes(preds, 0, 1).reshape(num_labels, -1).T
    target = np.swapaxes(target, 0, 1).reshape(num_labels, -1).T
    if not _in_unit_range(preds):
        preds = _sigmoid32(preds)
    ignored = (target == ignore_index) if ignore_index is not None else np.zeros(target.shape, dtype=bool)
    return preds, target, ignored


def multilabel_precision_recall_curve_update(preds, target, ignored, num_labels, thresholds) -> np.ndarray:
    """(T, L, 2, 2) histogram of key = above + 2*target + 4*label + 4*L*thr, ignored entries dropped
    (precision_recall_curve.py:769-793)."""
    thr = np.asarray(thresholds)
    T = len(thr)
    cm = np.zeros((T, num_labels, 2, 2), dtype=I64)
    t = np.asarray(target).astype(I64)
    for i in range(T):
        above = (preds >= thr[i])
        for tv in (0, 1):
            for pv in (0, 1):
                cm[i, :, tv, pv] = ((t == tv) & (above == bool(pv)) & ~ignored).sum(0)
    return cm


# --------------------------------------------------------------------------------------
# A9  precision_recall_curve.py:28-80   _binary_clf_curve (no sample weights)
# --------------------------------------------------------------------------------------
def binary_clf_curve(preds, target, pos_label=1):
    """Descending sort, keep the last index of every run of equal scores, cumulative positives there.
    tps is float32 in the reference (`target * 1.0` promotes long -> float32), fps = 1 + idx - tps."""
    preds = np.asarray(preds)
    target = np.asarray(target)
    order = np.argsort(-preds, kind="stable")
    p = preds[order]
    t = (target[order] == pos_label).astype(I64)
    distinct = np.nonzero(p[1:] - p[:-1])[0]
    idx = np.concatenate([distinct, np.array([t.size - 1], dtype=distinct.dtype)])
    tps = np.cumsum(t.astype(np.float32), dtype=np.float32)[idx]
    fps = (1 + idx).astype(np.float32) - tps
    return fps.astype(np.float32), tps, p[idx]


def binary_clf_curve_exact(preds, target, pos_label=1):
    """Same curve with exact int64 counts (what the CUDA path accumulates before the final cast)."""
    preds = np.asarray(preds)
    target = np.asarray(target)
    order = np.argsort(-preds, kind="stable")
    p = preds[order]
    t = (target[order] == pos_label).astype(I64)
    distinct = np.nonzero(p[1:] != p[:-1])[0]
    idx = np.concatenate([distinct, np.array([t.size - 1], dtype=distinct.dtype)])
    tps = np.cumsum(t, dtype=I64)[idx]
    fps = 1 + idx.astype(I64) - tps
    return fps, tps, p[idx]


# --------------------------------------------------------------------------------------
# A10  compute side: utilities/compute.py:46-55,88-96; roc.py:40-80; auroc.py:82-106;
#      precision_recall_curve.py:253-283; average_precision.py:70-75
# --------------------------------------------------------------------------------------
def safe_divide(num, denom):
    num = np.asarray(num)
    denom = np.asarray(denom).copy()
    denom[denom == 0] = 1
    return num.astype(np.float32) / denom.astype(np.float32)


def trapz(y, x, dtype=np.float32):
    y = np.asarray(y, dtype=dtype)
    x = np.asarray(x, dtype=dtype)
    return ((x[1:] - x[:-1]) * (y[1:] + y[:-1]) / dtype(2)).sum(dtype=dtype)


def binary_roc_compute_binned(confmat, thresholds):
    tps, fps, fns, tns = confmat[:, 1, 1], confmat[:, 0, 1], confmat[:, 1, 0], confmat[:, 0, 0]
    tpr = safe_divide(tps, tps + fns)[::-1]
    fpr = safe_divide(fps, fps + tns)[::-1]
    return fpr, tpr, np.asarray(thresholds)[::-1]


def binary_roc_compute_exact(preds, target, pos_label=1, dtype=np.float32):
    if dtype == np.float32:
        fps, tps, thr = binary_clf_curve(preds, target, pos_label)
    else:
        fps, tps, thr = binary_clf_curve_exact(preds, target, pos_label)
        fps, tps = fps.astype(dtype), tps.astype(dtype)
    tps = np.concatenate([np.zeros(1, dtype), tps])
    fps = np.concatenate([np.zeros(1, dtype), fps])
    thr = np.concatenate([np.ones(1, thr.dtype), thr])
    fpr = np.zeros_like(fps) if fps[-1] <= 0 else fps / fps[-1]
    tpr = np.zeros_like(tps) if tps[-1] <= 0 else tps / tps[-1]
    return fpr, tpr, thr


def binary_auroc_binned(confmat, thresholds):
    fpr, tpr, _ = binary_roc_compute_binned(confmat, thresholds)
    return trapz(tpr, fpr)


def binary_auroc_exact(preds, target, dtype=np.float32):
    fpr, tpr, _ = binary_roc_compute_exact(preds, target, dtype=dtype)
    return trapz(tpr, fpr, dtype)


def binary_precision_recall_binned(confmat, thresholds):
    tps, fps, fns = confmat[:, 1, 1], confmat[:, 0, 1], confmat[:, 1, 0]
    precision = np.concatenate([safe_divide(tps, tps + fps), np.ones(1, np.float32)])
    recall = np.concatenate([safe_divide(tps, tps + fns), np.zeros(1, np.float32)])
    return precision, recall, np.asarray(thresholds)


def binary_precision_recall_exact(preds, target, pos_label=1, dtype=np.float32):
    if dtype == np.float32:
        fps, tps, thr = binary_clf_curve(preds, target, pos_label)
    else:
        fps, tps, thr = binary_clf_curve_exact(preds, target, pos_label)
        fps, tps = fps.astype(dtype), tps.astype(dtype)
    with np.errstate(invalid="ignore", divide="ignore"):
        precision = tps / (tps + fps)
        recall = tps / tps[-1]
    precision = np.concatenate([precision[::-1], np.ones(1, dtype)])
    recall = np.concatenate([recall[::-1], np.zeros(1, dtype)])
    return precision, recall, thr[::-1].copy()


def average_precision_from_curve(precision, recall):
    return -((recall[1:] - recall[:-1]) * precision[:-1]).sum(dtype=precision.dtype)


def binary_average_precision_binned(confmat, thresholds):
    p, r, _ = binary_precision_recall_binned(confmat, thresholds)
    return average_precision_from_curve(p, r)


def binary_average_precision_exact(preds, target, dtype=np.float32):
    p, r, _ = binary_precision_recall_exact(preds, target, dtype=dtype)
    return average_precision_from_curve(p, r)


# --------------------------------------------------------------------------------------
# A13/A14  metric.py:427-457, utilities/distributed.py:97-147   cross-rank merge semantics
# --------------------------------------------------------------------------------------
def gather_all_tensors(per_rank: list) -> list:
    """What every rank ends up with after the reference's pad -> all_gather -> trim: the list of every
    rank's tensor, unchanged (utilities/distributed.py:124-147)."""
    ndim = {np.asarray(a).ndim for a in per_rank}
    if len(ndim) != 1:
        raise ValueError("all ranks must contribute tensors with the same number of dimensions")
    return [np.asarray(a).copy() for a in per_rank]


def sync_state(per_rank: list, reduce_fx: str):
    """State after `_sync_dist` (metric.py:442-457): stack + reduce for tensor states, concatenate for cat."""
    g = gather_all_tensors(per_rank)
    if reduce_fx == "cat":
        return np.concatenate([np.atleast_1d(a) for a in g], axis=0)
    stacked = np.stack(g)
    return {"sum": stacked.sum(0), "mean": stacked.mean(0), "max": stacked.max(0), "min": stacked.min(0)}[reduce_fx]
