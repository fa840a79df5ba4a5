"""O(C) / O(T) reducers that turn count states into scores (reference: utilities/compute.py:46-157).

On CPU tensors these are the reference's torch ops.  For int64 count states on a CUDA device the stat-score family
(`_fused_stat_reduce`) runs the whole chain in one kernel launch (`tmb_stat_reduce`), which is what `compute()` costs.
"""
from typing import Optional, Tuple

import torch
from torch import Tensor


def _fused_stat_reduce(kind: str, tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str],
                       multidim_average: str = "global", multilabel: bool = False, beta: float = 1.0) -> Optional[Tensor]:
    """precision / recall / specificity / accuracy / fbeta of CUDA int64 states in one launch, or None when the states
    are not served by the kernel (CPU tensors, other dtypes): the caller then runs the reference's torch ops."""
    if not tp.is_cuda:
        return None
    from torchmetrics_b200 import _native as N

    if not N.stat_reduce_supported(tp, fp, tn, fn, average):
        return None
    if average in ("micro", "macro", "weighted") and tp.ndim not in (0, 1 if multidim_average == "global" else 2):
        return None
    code = {"precision": N.STAT_PRECISION, "recall": N.STAT_RECALL, "specificity": N.STAT_SPECIFICITY,
            "accuracy": N.STAT_ACCURACY, "fbeta": N.STAT_FBETA}[kind]
    return N.stat_reduce(code, tp, fp, tn, fn, average, multilabel, beta)


def _safe_divide(num: Tensor, denom: Tensor) -> Tensor:
    """num / denom with 0-denominators replaced by 1; integer inputs are promoted to float32 (compute.py:46-55)."""
    denom = torch.where(denom == 0, torch.ones_like(denom), denom)
    num = num if num.is_floating_point() else num.float()
    denom = denom if denom.is_floating_point() else denom.float()
    return num / denom


def _adjust_weights_safe_divide(
    score: Tensor, average: Optional[str], multilabel: bool, tp: Tensor, fp: Tensor, fn: Tensor
) -> Tensor:
    """macro / weighted averaging of per-class scores; classes that never occur get weight 0 (compute.py:58-69)."""
    if average is None or average == "none":
        return score
    if average == "weighted":
        weights = tp + fn
    else:
        weights = torch.ones_like(score)
        if not multilabel:
            weights = torch.where(tp + fp + fn == 0, torch.zeros_like(weights), weights)
    return _safe_divide(weights * score, weights.sum(-1, keepdim=True)).sum(-1)


def _auc_compute_without_check(x: Tensor, y: Tensor, direction: float, axis: int = -1) -> Tensor:
    """Trapezoidal area, x assumed monotone (compute.py:88-96)."""
    with torch.no_grad():
        return torch.trapz(y, x, dim=axis) * direction


def _auc_format_inputs(x: Tensor, y: Tensor) -> Tuple[Tensor, Tensor]:
    x = x.squeeze() if x.ndim > 1 else x
    y = y.squeeze() if y.ndim > 1 else y
    if x.ndim > 1 or y.ndim > 1:
        raise ValueError(f"Expected both `x` and `y` tensor to be 1d, but got tensors with dimension {x.ndim} and {y.ndim}")
    if x.numel() != y.numel():
        raise ValueError(f"Expected the same number of elements in `x` and `y` tensor but received {x.numel()} and {y.numel()}")
    return x, y


def _auc_compute(x: Tensor, y: Tensor, reorder: bool = False) -> Tensor:
    with torch.no_grad():
        if reorder:
            x, order = torch.sort(x, stable=True)
            y = y[order]
        dx = x[1:] - x[:-1]
        if (dx < 0).any():
            if not (dx <= 0).all():
                raise ValueError(
                    "The `x` tensor is neither increasing or decreasing. Try setting the reorder argument to `True`."
                )
            return _auc_compute_without_check(x, y, -1.0)
        return _auc_compute_without_check(x, y, 1.0)


def auc(x: Tensor, y: Tensor, reorder: bool = False) -> Tensor:
    x, y = _auc_format_inputs(x, y)
    return _auc_compute(x, y, reorder=reorder)


def interp(x: Tensor, xp: Tensor, fp: Tensor) -> Tensor:
    """Piecewise-linear interpolation of (xp, fp) at x, xp increasing (compute.py:134-157)."""
    slope = _safe_divide(fp[1:] - fp[:-1], xp[1:] - xp[:-1])
    offset = fp[:-1] - slope * xp[:-1]
    seg = torch.sum(torch.ge(x[:, None], xp[None, :]), 1) - 1
    seg = torch.clamp(seg, 0, len(slope) - 1)
    return slope[seg] * x + offset[seg]
