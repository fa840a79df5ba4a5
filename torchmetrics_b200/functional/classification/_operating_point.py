"""Operating points on a curve: recall@precision, precision@recall, sensitivity@specificity, specificity@sensitivity.

Reference: functional/classification/{recall_fixed_precision,precision_fixed_recall,sensitivity_specificity,
specificity_sensitivity}.py.  All four share one shape: build the binned or exact curve state exactly as
`*_precision_recall_curve` does (same K3 / K4 kernels), turn it into a curve with
`_*_precision_recall_curve_compute` or `_*_roc_compute`, and pick one point per class with a selection rule.  Only the
O(T) selection rules are specific to this file; the four public modules are instantiations of `make_family`.
"""
from typing import Callable, List, Optional, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.functional.classification import precision_recall_curve as P
from torchmetrics_b200.functional.classification import roc as R
from torchmetrics_b200.utilities.enums import ClassificationTask

_Thr = Optional[Union[int, List[float], Tensor]]


def _lexargmax(x: Tensor) -> Tensor:
    """Row indices of the lexicographic maximum of the columns of `x` (recall_fixed_precision.py:45-67)."""
    rows = torch.arange(x.shape[0], device=x.device)
    for k in range(x.shape[1]):
        col = x[rows, k]
        rows = rows[col == col.max()]
        if rows.numel() < 2:
            break
    return rows


def _best_row(first: Tensor, second: Tensor, thresholds: Tensor, keep: Tensor) -> Tuple[Tensor, Tensor]:
    """Among the curve points with `keep`, the lexicographic maximum of (first, second, threshold); (0, 1e6) when none
    qualifies or the best `first` is 0."""
    n = min(first.shape[0], second.shape[0], thresholds.shape[0])
    table = torch.stack([first[:n], second[:n], thresholds[:n].to(first.dtype)], dim=1)[keep[:n]]
    best = torch.tensor(0.0, device=first.device, dtype=first.dtype)
    at = torch.tensor(0)
    if table.shape[0] > 0:
        row = table[_lexargmax(table)[0]]
        best, at = row[0], row[2]
    if best == 0.0:
        at = torch.tensor(1e6, device=thresholds.device, dtype=thresholds.dtype)
    return best, at


def _recall_at_precision(precision: Tensor, recall: Tensor, thresholds: Tensor, min_precision: float):
    """recall_fixed_precision.py:70-89: highest recall among points with precision >= min_precision."""
    return _best_row(recall, precision, thresholds, precision >= min_precision)


def _precision_at_recall(precision: Tensor, recall: Tensor, thresholds: Tensor, min_recall: float):
    """precision_fixed_recall.py:42-58: `max((p, r, t) ...)` over points with recall >= min_recall, vectorised."""
    return _best_row(precision, recall, thresholds, recall >= min_recall)


def _first_argmax_where(value: Tensor, constraint: Tensor, thresholds: Tensor, minimum: float):
    """sensitivity_specificity.py:47-63 / specificity_sensitivity.py:47-63: first maximum of `value` among the points
    whose `constraint` reaches `minimum`; (0, 1e6) when there is none."""
    keep = constraint >= minimum
    if not bool(keep.any()):
        return (torch.tensor(0.0, device=value.device, dtype=value.dtype),
                torch.tensor(1e6, device=thresholds.device, dtype=thresholds.dtype))
    value, thresholds = value[keep], thresholds[keep]
    i = torch.argmax(value)
    return value[i], thresholds[i]


def _sensitivity_at_specificity(sensitivity: Tensor, specificity: Tensor, thresholds: Tensor, min_specificity: float):
    return _first_argmax_where(sensitivity, specificity, thresholds, min_specificity)


def _specificity_at_sensitivity(specificity: Tensor, sensitivity: Tensor, thresholds: Tensor, min_sensitivity: float):
    return _first_argmax_where(specificity, sensitivity, thresholds, min_sensitivity)


def _convert_fpr_to_specificity(fpr: Tensor) -> Tensor:
    return 1 - fpr


def _check_minimum(name: str, value: float) -> None:
    # the reference's condition (`not isinstance(float) and not in range`) is kept as written
    if not isinstance(value, float) and not (0 <= value <= 1):
        raise ValueError(f"Expected argument `{name}` to be an float in the [0,1] range, but got {value}")


class Family:
    """The three tasks + dispatcher of one operating-point metric."""

    def __init__(self, arg_name: str, curve: str, select: Callable) -> None:
        self.arg_name = arg_name  # name of the minimum in error messages
        self.curve = curve        # "pr": select(precision, recall, thr, m); "roc": select(a, b, thr, m) on tpr / 1-fpr
        self.select = select

    # ---- curve -> per-class points
    def _pairs(self, c1, c2):
        if self.curve == "pr":
            return c1, c2  # precision, recall
        # roc compute returns (fpr, tpr): the selection rules take (value, constraint)
        spec = [_convert_fpr_to_specificity(f) for f in c1] if isinstance(c1, (list, tuple)) or c1.ndim > 1 \
            else _convert_fpr_to_specificity(c1)
        return (c2, spec) if self.select is _sensitivity_at_specificity else (spec, c2)

    def binary_compute(self, state, thresholds, minimum: float, pos_label: int = 1, reduce_fn: Optional[Callable] = None):
        fn = reduce_fn or self.select
        if self.curve == "pr":
            a, b, thr = P._binary_precision_recall_curve_compute(state, thresholds, pos_label)
        else:
            fpr, tpr, thr = R._binary_roc_compute(state, thresholds, pos_label)
            a, b = self._pairs(fpr, tpr)
        return fn(a, b, thr, minimum)

    def _per_class(self, a, b, thr, state, minimum: float, fn: Callable):
        if isinstance(state, Tensor):
            res = [fn(x, y, thr, minimum) for x, y in zip(a, b)]
        else:
            res = [fn(x, y, t, minimum) for x, y, t in zip(a, b, thr)]
        return torch.stack([r[0] for r in res]), torch.stack([r[1] for r in res])

    def multiclass_compute(self, state, num_classes: int, thresholds, minimum: float,
                           reduce_fn: Optional[Callable] = None):
        fn = reduce_fn or self.select
        if self.curve == "pr":
            a, b, thr = P._multiclass_precision_recall_curve_compute(state, num_classes, thresholds)
        else:
            fpr, tpr, thr = R._multiclass_roc_compute(state, num_classes, thresholds)
            a, b = self._pairs(fpr, tpr)
        return self._per_class(a, b, thr, state, minimum, fn)

    def multilabel_compute(self, state, num_labels: int, thresholds, ignore_index: Optional[int], minimum: float,
                           reduce_fn: Optional[Callable] = None):
        fn = reduce_fn or self.select
        if self.curve == "pr":
            a, b, thr = P._multilabel_precision_recall_curve_compute(state, num_labels, thresholds, ignore_index)
        else:
            fpr, tpr, thr = R._multilabel_roc_compute(state, num_labels, thresholds, ignore_index)
            a, b = self._pairs(fpr, tpr)
        return self._per_class(a, b, thr, state, minimum, fn)

    # ---- argument validation
    def binary_arg_validation(self, minimum: float, thresholds: _Thr = None, ignore_index: Optional[int] = None):
        P._binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
        _check_minimum(self.arg_name, minimum)

    def multiclass_arg_validation(self, num_classes: int, minimum: float, thresholds: _Thr = None,
                                  ignore_index: Optional[int] = None):
        P._multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index)
        _check_minimum(self.arg_name, minimum)

    def multilabel_arg_validation(self, num_labels: int, minimum: float, thresholds: _Thr = None,
                                  ignore_index: Optional[int] = None):
        P._multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
        _check_minimum(self.arg_name, minimum)

    # ---- one-shot functionals: the state comes from the same fused K3 / K4 path as the curve functionals
    def binary(self, preds: Tensor, target: Tensor, minimum: float, thresholds: _Thr = None,
               ignore_index: Optional[int] = None, validate_args: bool = True):
        if validate_args:
            self.binary_arg_validation(minimum, thresholds, ignore_index)
        state, thr = P._binary_curve_state(preds, target, thresholds, ignore_index, validate=validate_args)
        return self.binary_compute(state, thr, minimum)

    def multiclass(self, preds: Tensor, target: Tensor, num_classes: int, minimum: float, thresholds: _Thr = None,
                   ignore_index: Optional[int] = None, validate_args: bool = True):
        if validate_args:
            self.multiclass_arg_validation(num_classes, minimum, thresholds, ignore_index)
        state, thr = P._multiclass_curve_state(preds, target, num_classes, thresholds, ignore_index,
                                               validate=validate_args)
        return self.multiclass_compute(state, num_classes, thr, minimum)

    def multilabel(self, preds: Tensor, target: Tensor, num_labels: int, minimum: float, thresholds: _Thr = None,
                   ignore_index: Optional[int] = None, validate_args: bool = True):
        if validate_args:
            self.multilabel_arg_validation(num_labels, minimum, thresholds, ignore_index)
        state, thr = P._multilabel_curve_state(preds, target, num_labels, thresholds, ignore_index,
                                               validate=validate_args)
        return self.multilabel_compute(state, num_labels, thr, ignore_index, minimum)

    def dispatch(self, preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
                 minimum: float, thresholds: _Thr = None, num_classes: Optional[int] = None,
                 num_labels: Optional[int] = None, ignore_index: Optional[int] = None, validate_args: bool = True):
        task = ClassificationTask.from_str(task)
        if task == ClassificationTask.BINARY:
            return self.binary(preds, target, minimum, thresholds, ignore_index, validate_args)
        if task == ClassificationTask.MULTICLASS:
            if not isinstance(num_classes, int):
                raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
            return self.multiclass(preds, target, num_classes, minimum, thresholds, ignore_index, validate_args)
        if task == ClassificationTask.MULTILABEL:
            if not isinstance(num_labels, int):
                raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
            return self.multilabel(preds, target, num_labels, minimum, thresholds, ignore_index, validate_args)
        raise ValueError(f"Not handled value: {task}")


RECALL_AT_PRECISION = Family("min_precision", "pr", _recall_at_precision)
PRECISION_AT_RECALL = Family("min_precision", "pr", _precision_at_recall)
SENSITIVITY_AT_SPECIFICITY = Family("min_specificity", "roc", _sensitivity_at_specificity)
SPECIFICITY_AT_SENSITIVITY = Family("min_sensitivity", "roc", _specificity_at_sensitivity)
