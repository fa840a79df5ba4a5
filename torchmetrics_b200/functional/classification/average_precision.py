"""Average precision (reference: functional/classification/average_precision.py)."""
from importlib import import_module
from typing import List, Optional, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

P = import_module("torchmetrics_b200.functional.classification.precision_recall_curve")
from torchmetrics_b200.functional.classification._exact import fused_binary_average_precision
from torchmetrics_b200.utilities.compute import _safe_divide
from torchmetrics_b200.utilities.data import _bincount
from torchmetrics_b200.utilities.enums import ClassificationTask
from torchmetrics_b200.utilities.prints import rank_zero_warn


def _reduce_average_precision(precision: Union[Tensor, List[Tensor]], recall: Union[Tensor, List[Tensor]],
                              average: Optional[Literal["macro", "weighted", "none"]] = "macro",
                              weights: Optional[Tensor] = None) -> Tensor:
    """Step-function area -sum((r[i+1]-r[i]) * p[i]) per class, then averaging (average_precision.py:43-67)."""
    if isinstance(precision, Tensor) and isinstance(recall, Tensor):
        res = -torch.sum((recall[:, 1:] - recall[:, :-1]) * precision[:, :-1], 1)
    else:
        res = torch.stack([-torch.sum((r[1:] - r[:-1]) * p[:-1]) for p, r in zip(precision, recall)])
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


def _binary_average_precision_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], thresholds: Optional[Tensor]) -> Tensor:
    if thresholds is None:
        fused = fused_binary_average_precision(state)  # exact mode: reduced next to the sort (K4b)
        if fused is not None:
            return fused
    precision, recall, _ = P._binary_precision_recall_curve_compute(state, thresholds)
    return -torch.sum((recall[1:] - recall[:-1]) * precision[:-1])


def binary_average_precision(preds: Tensor, target: Tensor,
                             thresholds: Optional[Union[int, List[float], Tensor]] = None,
                             ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    if validate_args:
        P._binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
    state, thr = P._binary_curve_state(preds, target, thresholds, ignore_index, validate=validate_args)
    return _binary_average_precision_compute(state, thr)


def _multiclass_average_precision_arg_validation(num_classes: int, average: Optional[str] = "macro",
                                                 thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                                 ignore_index: Optional[int] = None) -> None:
    P._multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index)
    allowed = ("macro", "weighted", "none", None)
    if average not in allowed:
        raise ValueError(f"Expected argument `average` to be one of {allowed} but got {average}")


def _multiclass_average_precision_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], num_classes: int,
                                          average: Optional[str] = "macro", thresholds: Optional[Tensor] = None) -> Tensor:
    precision, recall, _ = P._multiclass_precision_recall_curve_compute(state, num_classes, thresholds)
    weights = (_bincount(state[1], minlength=num_classes).float() if thresholds is None
               else state[0][:, 1, :].sum(-1))
    return _reduce_average_precision(precision, recall, average, weights=weights)


def multiclass_average_precision(preds: Tensor, target: Tensor, num_classes: int,
                                 average: Optional[Literal["macro", "weighted", "none"]] = "macro",
                                 thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                 ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    if validate_args:
        _multiclass_average_precision_arg_validation(num_classes, average, thresholds, ignore_index)
    state, thr = P._multiclass_curve_state(preds, target, num_classes, thresholds, ignore_index,
                                           validate=validate_args)
    return _multiclass_average_precision_compute(state, num_classes, average, thr)


def _multilabel_average_precision_arg_validation(num_labels: int, average: Optional[str] = "macro",
                                                 thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                                 ignore_index: Optional[int] = None) -> None:
    P._multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
    allowed = ("micro", "macro", "weighted", "none", None)
    if average not in allowed:
        raise ValueError(f"Expected argument `average` to be one of {allowed} but got {average}")


def _multilabel_average_precision_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], num_labels: int,
                                          average: Optional[str], thresholds: Optional[Tensor],
                                          ignore_index: Optional[int] = None) -> Tensor:
    if average == "micro":
        if isinstance(state, Tensor) and thresholds is not None:
            state = state.sum(1)
        else:
            preds, target = state[0].flatten(), state[1].flatten()
            if ignore_index is not None:
                keep = target != ignore_index
                preds, target = preds[keep], target[keep]
            state = (preds, target)
        return _binary_average_precision_compute(state, thresholds)
    precision, recall, _ = P._multilabel_precision_recall_curve_compute(state, num_labels, thresholds, ignore_index)
    weights = (state[1] == 1).sum(dim=0).float() if thresholds is None else state[0][:, 1, :].sum(-1)
    return _reduce_average_precision(precision, recall, average, weights=weights)


def multilabel_average_precision(preds: Tensor, target: Tensor, num_labels: int,
                                 average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                                 thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                 ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    if validate_args:
        _multilabel_average_precision_arg_validation(num_labels, average, thresholds, ignore_index)
    state, thr = P._multilabel_curve_state(preds, target, num_labels, thresholds, ignore_index,
                                           validate=validate_args)
    return _multilabel_average_precision_compute(state, num_labels, average, thr, ignore_index)


def average_precision(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
                      thresholds: Optional[Union[int, List[float], Tensor]] = None, num_classes: Optional[int] = None,
                      num_labels: Optional[int] = None,
                      average: Optional[Literal["macro", "weighted", "none"]] = "macro",
                      ignore_index: Optional[int] = None, validate_args: bool = True) -> Optional[Tensor]:
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_average_precision(preds, target, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_average_precision(preds, target, num_classes, average, thresholds, ignore_index,
                                            validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_average_precision(preds, target, num_labels, average, thresholds, ignore_index,
                                            validate_args)
    return None
