"""AUROC classes (reference: classification/auroc.py)."""
from importlib import import_module
from typing import Any, Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_b200.classification.precision_recall_curve import (
    BinaryPrecisionRecallCurve,
    MulticlassPrecisionRecallCurve,
    MultilabelPrecisionRecallCurve,
    _curve_task,
    _Thresholds,
)
F = import_module("torchmetrics_b200.functional.classification.auroc")
from torchmetrics_b200.metric import Metric


class BinaryAUROC(BinaryPrecisionRecallCurve):
    is_differentiable = False
    higher_is_better = True
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0

    def __init__(self, max_fpr: Optional[float] = None, thresholds: _Thresholds = None,
                 ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(thresholds=thresholds, ignore_index=ignore_index, validate_args=False, **kwargs)
        if validate_args:
            F._binary_auroc_arg_validation(max_fpr, thresholds, ignore_index)
        self.validate_args = validate_args
        self.max_fpr = max_fpr

    _supports_partition = True

    def compute(self) -> Tensor:
        sums = self._partitioned_sums()
        if sums is not None:
            if self.max_fpr not in (None, 1):
                raise ValueError("`distributed_exact='partition'` does not support `max_fpr`")
            area2, _, npos, n = sums
            nneg = n - npos
            if npos <= 0 or nneg <= 0:  # the reference warns and integrates an all-zero curve (roc.py:60-75)
                return torch.tensor(0.0, device=self.device)
            return torch.tensor(area2 / (2.0 * npos * nneg), dtype=torch.float32, device=self.device)
        return F._binary_auroc_compute(self._state(), self.thresholds, self.max_fpr)


class MulticlassAUROC(MulticlassPrecisionRecallCurve):
    is_differentiable = False
    higher_is_better = True
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0
    plot_legend_name = "Class"

    def __init__(self, num_classes: int, average: Optional[Literal["macro", "weighted", "none"]] = "macro",
                 thresholds: _Thresholds = None, ignore_index: Optional[int] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(num_classes=num_classes, thresholds=thresholds, ignore_index=ignore_index,
                         validate_args=False, **kwargs)
        if validate_args:
            F._multiclass_auroc_arg_validation(num_classes, average, thresholds, ignore_index)
        self.average = average
        self.validate_args = validate_args

    def update(self, preds: Tensor, target: Tensor) -> None:
        # the state is always per class; `average` only matters in compute
        avg, self.average = self.average, None
        try:
            super().update(preds, target)
        finally:
            self.average = avg

    def compute(self) -> Tensor:
        return F._multiclass_auroc_compute(self._state(), self.num_classes, self.average, self.thresholds)


class MultilabelAUROC(MultilabelPrecisionRecallCurve):
    is_differentiable = False
    higher_is_better = True
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0
    plot_legend_name = "Label"

    def __init__(self, num_labels: int, average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                 thresholds: _Thresholds = None, ignore_index: Optional[int] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(num_labels=num_labels, thresholds=thresholds, ignore_index=ignore_index, validate_args=False,
                         **kwargs)
        if validate_args:
            F._multilabel_auroc_arg_validation(num_labels, average, thresholds, ignore_index)
        self.average = average
        self.validate_args = validate_args

    def compute(self) -> Tensor:
        return F._multilabel_auroc_compute(self._state(), self.num_labels, self.average, self.thresholds,
                                           self.ignore_index)


class AUROC(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], thresholds: _Thresholds = None,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["macro", "weighted", "none"]] = "macro", max_fpr: Optional[float] = None,
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        return _curve_task(task, BinaryAUROC, MulticlassAUROC, MultilabelAUROC, thresholds, num_classes, num_labels,
                           ignore_index, validate_args, kwargs, multiclass_extra={"average": average},
                           multilabel_extra={"average": average}, binary_extra={"max_fpr": max_fpr})
