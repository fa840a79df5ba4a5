"""ROC classes (reference: classification/roc.py): same states as the PR-curve classes, different `compute`."""
from importlib import import_module
from typing import Any, Optional

from typing_extensions import Literal

from torchmetrics_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_b200.classification.precision_recall_curve import (
    BinaryPrecisionRecallCurve,
    MulticlassPrecisionRecallCurve,
    MultilabelPrecisionRecallCurve,
    _curve_task,
    _Thresholds,
)
F = import_module("torchmetrics_b200.functional.classification.roc")
from torchmetrics_b200.metric import Metric


class BinaryROC(BinaryPrecisionRecallCurve):
    def compute(self):
        return F._binary_roc_compute(self._state(), self.thresholds)


class MulticlassROC(MulticlassPrecisionRecallCurve):
    def compute(self):
        return F._multiclass_roc_compute(self._state(), self.num_classes, self.thresholds, self.average)


class MultilabelROC(MultilabelPrecisionRecallCurve):
    def compute(self):
        return F._multilabel_roc_compute(self._state(), self.num_labels, self.thresholds, self.ignore_index)


class ROC(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], thresholds: _Thresholds = None,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        return _curve_task(task, BinaryROC, MulticlassROC, MultilabelROC, thresholds, num_classes, num_labels,
                           ignore_index, validate_args, kwargs)
