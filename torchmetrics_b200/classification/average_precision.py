"""AveragePrecision classes (reference: classification/average_precision.py)."""
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
F = import_module("torchmetrics_b200.functional.classification.average_precision")
from torchmetrics_b200.metric import Metric


class BinaryAveragePrecision(BinaryPrecisionRecallCurve):
    is_differentiable = False
    higher_is_better = True
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0

    _supports_partition = True

    def compute(self) -> Tensor:
        sums = self._partitioned_sums()
        if sums is not None:
            _, apsum, npos, _ = sums
            if npos <= 0:  # no positive sample anywhere: recall is 0/0 (average_precision.py:39-46 gives nan)
                return torch.tensor(float("nan"), device=self.device)
            return torch.tensor(apsum / npos, dtype=torch.float32, device=self.device)
        return F._binary_average_precision_compute(self._state(), self.thresholds)


class MulticlassAveragePrecision(MulticlassPrecisionRecallCurve):
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
            F._multiclass_average_precision_arg_validation(num_classes, average, thresholds, ignore_index)
        self.average = average
        self.validate_args = validate_args

    def update(self, preds: Tensor, target: Tensor) -> None:
        avg, self.average = self.average, None
        try:
            super().update(preds, target)
        finally:
            self.average = avg

    def compute(self) -> Tensor:
        return F._multiclass_average_precision_compute(self._state(), self.num_classes, self.average, self.thresholds)


class MultilabelAveragePrecision(MultilabelPrecisionRecallCurve):
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
            F._multilabel_average_precision_arg_validation(num_labels, average, thresholds, ignore_index)
        self.average = average
        self.validate_args = validate_args

    def compute(self) -> Tensor:
        return F._multilabel_average_precision_compute(self._state(), self.num_labels, self.average, self.thresholds,
                                                       self.ignore_index)


class AveragePrecision(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], thresholds: _Thresholds = None,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["macro", "weighted", "none"]] = "macro",
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        return _curve_task(task, BinaryAveragePrecision, MulticlassAveragePrecision, MultilabelAveragePrecision,
                           thresholds, num_classes, num_labels, ignore_index, validate_args, kwargs,
                           multiclass_extra={"average": average}, multilabel_extra={"average": average})
