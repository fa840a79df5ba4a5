"""Area under the ROC curve (reference: functional/classification/auroc.py)."""
from importlib import import_module
from typing import List, Optional, Tuple, Union

import torch
from torch import Tensor, tensor
from typing_extensions import Literal

P = import_module("torchmetrics_b200.functional.classification.precision_recall_curve")
from torchmetrics_b200.functional.classification._exact import fused_binary_auroc
from torchmetrics_b200.functional.classification.roc import (
    _binary_roc_compute,
    _multiclass_roc_compute,
    _multilabel_roc_compute,
)
from torchmetrics_b200.utilities.compute import _auc_compute_without_check, _safe_divide
from torchmetrics_b200.utilities.data import _bincount
from torchmetrics_b200.utilities.enums import ClassificationTask
from torchmetrics_b200.utilities.prints import rank_zero_warn


def _reduce_auroc(fpr: Union[Tensor, List[Tensor]], tpr: Union[Tensor, List[Tensor]],
                  average: Optional[Literal["macro", "weighted", "none"]] = "macro",
                  weights: Optional[Tensor] = None) -> Tensor:
    """Per-class trapezoid areas, then macro / weighted mean over the non-NaN ones (auroc.py:45-69)."""
    if isinstance(fpr, Tensor) and isinstance(tpr, Tensor):
        res = _auc_compute_without_check(fpr, tpr, 1.0, axis=1)
    else:
        res = torch.stack([_auc_compute_without_check(x, y, 1.0) for x, y in zip(fpr, tpr)])
    if average is None or average == "none":
        return res
    if torch.isnan(res).any():
        rank_zero_warn(
            f"Average precision score for one or more classes was `nan`. Ignoring these classes in {average}-average",
            UserWarning,
        )
    ok = ~torch.isnan(res)
    if average == "macro":
        return res[ok].mean()
    if average == "weighted" and weights is not None:
        w = _safe_divide(weights[ok], weights[ok].sum())
        return (res[ok] * w).sum()
    raise ValueError("Received an incompatible combinations of inputs to make reduction.")


def _binary_auroc_arg_validation(max_fpr: Optional[float] = None,
                                 thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                 ignore_index: Optional[int] = None) -> None:
    P._binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
    if max_fpr is not None and not isinstance(max_fpr, float) and 0 < max_fpr <= 1:
        raise ValueError(f"Arguments `max_fpr` should be a float in range (0, 1], but got: {max_fpr}")


def _binary_auroc_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], thresholds: Optional[Tensor],
                          max_fpr: Optional[float] = None, pos_label: int = 1) -> Tensor:
    """Full or partial (McClish-corrected) AUROC (auroc.py:82-106).  Exact mode without `max_fpr` never builds the
    curve: the trapezoid sum is reduced next to the sort (K4b)."""
    if thresholds is None and (max_fpr is None or max_fpr == 1):
        fused = fused_binary_auroc(state, pos_label)
        if fused is not None:
            return fused
    fpr, tpr, _ = _binary_roc_compute(state, thresholds, pos_label)
    if max_fpr is None or max_fpr == 1 or fpr.sum() == 0 or tpr.sum() == 0:
        return _auc_compute_without_check(fpr, tpr, 1.0)
    max_area: Tensor = tensor(max_fpr, device=fpr.device)
    stop = torch.bucketize(max_area, fpr, out_int32=True, right=True)
    weight = (max_area - fpr[stop - 1]) / (fpr[stop] - fpr[stop - 1])
    interp_tpr: Tensor = torch.lerp(tpr[stop - 1], tpr[stop], weight)
    tpr = torch.cat([tpr[:stop], interp_tpr.view(1)])
    fpr = torch.cat([fpr[:stop], max_area.view(1)])
    partial_auc = _auc_compute_without_check(fpr, tpr, 1.0)
    min_area: Tensor = 0.5 * max_area ** 2
    return 0.5 * (1 + (partial_auc - min_area) / (max_area - min_area))


def binary_auroc(preds: Tensor, target: Tensor, max_fpr: Optional[float] = None,
                 thresholds: Optional[Union[int, List[float], Tensor]] = None, ignore_index: Optional[int] = None,
                 validate_args: bool = True) -> Tensor:
    if validate_args:
        _binary_auroc_arg_validation(max_fpr, thresholds, ignore_index)
    state, thr = P._binary_curve_state(preds, target, thresholds, ignore_index, validate=validate_args)
    return _binary_auroc_compute(state, thr, max_fpr)


def _multiclass_auroc_arg_validation(num_classes: int, average: Optional[str] = "macro",
                                     thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                     ignore_index: Optional[int] = None) -> None:
    P._multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index)
    allowed = ("macro", "weighted", "none", None)
    if average not in allowed:
        raise ValueError(f"Expected argument `average` to be one of {allowed} but got {average}")


def _multiclass_auroc_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], num_classes: int,
                              average: Optional[str] = "macro", thresholds: Optional[Tensor] = None) -> Tensor:
    fpr, tpr, _ = _multiclass_roc_compute(state, num_classes, thresholds)
    weights = (_bincount(state[1], minlength=num_classes).float() if thresholds is None
               else state[0][:, 1, :].sum(-1))
    return _reduce_auroc(fpr, tpr, average, weights=weights)


def multiclass_auroc(preds: Tensor, target: Tensor, num_classes: int,
                     average: Optional[Literal["macro", "weighted", "none"]] = "macro",
                     thresholds: Optional[Union[int, List[float], Tensor]] = None, ignore_index: Optional[int] = None,
                     validate_args: bool = True) -> Tensor:
    if validate_args:
        _multiclass_auroc_arg_validation(num_classes, average, thresholds, ignore_index)
    state, thr = P._multiclass_curve_state(preds, target, num_classes, thresholds, ignore_index,
                                           validate=validate_args)
    return _multiclass_auroc_compute(state, num_classes, average, thr)


def _multilabel_auroc_arg_validation(num_labels: int, average: Optional[str],
                                     thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                     ignore_index: Optional[int] = None) -> None:
    P._multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
    allowed = ("micro", "macro", "weighted", "none", None)
    if average not in allowed:
        raise ValueError(f"Expected argument `average` to be one of {allowed} but got {average}")


def _multilabel_auroc_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], num_labels: int, average: Optional[str],
                              thresholds: Optional[Tensor], ignore_index: Optional[int] = None) -> Tensor:
    if average == "micro":
        if isinstance(state, Tensor) and thresholds is not None:
            return _binary_auroc_compute(state.sum(1), thresholds, max_fpr=None)
        preds, target = state[0].flatten(), state[1].flatten()
        if ignore_index is not None:
            keep = target != ignore_index
            preds, target = preds[keep], target[keep]
        return _binary_auroc_compute((preds, target), thresholds, max_fpr=None)
    fpr, tpr, _ = _multilabel_roc_compute(state, num_labels, thresholds, ignore_index)
    weights = (state[1] == 1).sum(dim=0).float() if thresholds is None else state[0][:, 1, :].sum(-1)
    return _reduce_auroc(fpr, tpr, average, weights=weights)


def multilabel_auroc(preds: Tensor, target: Tensor, num_labels: int,
                     average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                     thresholds: Optional[Union[int, List[float], Tensor]] = None, ignore_index: Optional[int] = None,
                     validate_args: bool = True) -> Tensor:
    if validate_args:
        _multilabel_auroc_arg_validation(num_labels, average, thresholds, ignore_index)
    state, thr = P._multilabel_curve_state(preds, target, num_labels, thresholds, ignore_index,
                                           validate=validate_args)
    return _multilabel_auroc_compute(state, num_labels, average, thr, ignore_index)


def auroc(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
          thresholds: Optional[Union[int, List[float], Tensor]] = None, num_classes: Optional[int] = None,
          num_labels: Optional[int] = None, average: Optional[Literal["macro", "weighted", "none"]] = "macro",
          max_fpr: Optional[float] = None, ignore_index: Optional[int] = None, validate_args: bool = True
          ) -> Optional[Tensor]:
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_auroc(preds, target, max_fpr, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_auroc(preds, target, num_classes, average, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_auroc(preds, target, num_labels, average, thresholds, ignore_index, validate_args)
    return None
