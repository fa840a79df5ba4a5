"""Exact (sort-based) binary curve: descending sort of the scores, cumulative positives at the last index of every
run of equal scores (reference: functional/classification/precision_recall_curve.py:28-80)."""
from typing import Optional, Tuple

import torch
from torch import Tensor

from torchmetrics_b200 import _native


def exact_clf_curve(preds: Tensor, target: Tensor, pos_label: int = 1) -> Tuple[Tensor, Tensor, Tensor]:
    """(fps, tps, thresholds) with exact int64 counts; thresholds are the distinct scores, descending."""
    _native.require_cuda(preds, target)
    return _native.exact_curve(preds, target, pos_label)


def _fusable(state) -> bool:
    return (isinstance(state, (tuple, list)) and len(state) == 2 and isinstance(state[0], Tensor)
            and state[0].is_cuda and state[1].is_cuda and state[0].is_floating_point() and state[0].numel() > 0)


def fused_binary_auroc(state, pos_label: int = 1) -> Optional[Tensor]:
    """Exact AUROC of (preds, target) reduced on the device, or None when the caller should take the curve path
    (degenerate targets, where the reference warns and returns zeros)."""
    if not _fusable(state):
        return None
    preds, target = state
    area2, _, npos, _ = _native.exact_auroc_ap(preds, target, pos_label)
    nneg = preds.numel() - npos
    if npos <= 0 or nneg <= 0:
        return None
    return torch.tensor(area2 / (2.0 * npos * nneg), dtype=torch.float32, device=preds.device)


def fused_binary_average_precision(state, pos_label: int = 1) -> Optional[Tensor]:
    """Exact average precision reduced on the device, or None when there is no positive (reference: nan path)."""
    if not _fusable(state):
        return None
    preds, target = state
    _, apsum, npos, _ = _native.exact_auroc_ap(preds, target, pos_label)
    if npos <= 0:
        return None
    return torch.tensor(apsum / npos, dtype=torch.float32, device=preds.device)
