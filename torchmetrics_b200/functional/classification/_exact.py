"""Exact (sort-based) binary curve: descending sort of the scores, cumulative positives at the last index of every
run of equal scores (reference: functional/classification/precision_recall_curve.py:28-80)."""
from typing import Tuple

import torch
from torch import Tensor

from torchmetrics_b200 import _native


def exact_clf_curve(preds: Tensor, target: Tensor, pos_label: int = 1) -> Tuple[Tensor, Tensor, Tensor]:
    """(fps, tps, thresholds) with exact int64 counts; thresholds are the distinct scores, descending."""
    _native.require_cuda(preds, target)
    return _native.exact_curve(preds, target, pos_label)
