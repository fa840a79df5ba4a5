"""Stateful confusion matrices (reference: classification/confusion_matrix.py:51-534)."""
from importlib import import_module
from typing import Any, Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200 import _native
from torchmetrics_b200.classification.base import _ClassificationTaskWrapper
F = import_module("torchmetrics_b200.functional.classification.confusion_matrix")
C = import_module("torchmetrics_b200.functional.classification._checks")
from torchmetrics_b200.metric import Metric
from torchmetrics_b200.utilities.enums import ClassificationTask


class _ConfmatBase(Metric):
    is_differentiable = False
    higher_is_better = None
    full_state_update = False
    confmat: Tensor

    def _check_device_status(self, num_classes: Optional[int] = None) -> None:
        # label-range errors are flagged on the device during update and surfaced here, at the first sync point
        if self.confmat.is_cuda:
            _native.raise_on_status(self.confmat.device, num_classes)


class BinaryConfusionMatrix(_ConfmatBase):
    """2x2 confusion matrix, state `confmat` (sum-reduced). Rows = target, columns = prediction."""

    def __init__(self, threshold: float = 0.5, ignore_index: Optional[int] = None,
                 normalize: Optional[Literal["true", "pred", "all", "none"]] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(**kwargs)
        if validate_args:
            F._binary_confusion_matrix_arg_validation(threshold, ignore_index, normalize)
        self.threshold = threshold
        self.ignore_index = ignore_index
        self.normalize = normalize
        self.validate_args = validate_args
        self.add_state("confmat", torch.zeros(2, 2, dtype=torch.long), dist_reduce_fx="sum")

    def update(self, preds: Tensor, target: Tensor) -> None:
        fv = C.FusedValidation(self.validate_args, preds, target)
        if self.validate_args:
            with fv:
                F._binary_confusion_matrix_tensor_validation(preds, target, self.ignore_index)

        def step(out: Tensor) -> None:
            F._binary_confusion_matrix_update_fused(preds, target, self.threshold, self.ignore_index, out=out)

        step(self.confmat)
        fv.finish(preds.device, 2, undo=lambda: C.rollback(self.confmat, step), owner=self)

    def compute(self) -> Tensor:
        self._check_device_status(2)
        return F._binary_confusion_matrix_compute(self.confmat, self.normalize)


class MulticlassConfusionMatrix(_ConfmatBase):
    """CxC confusion matrix; one K1 pass per update straight into the int64 state."""

    def __init__(self, num_classes: int, ignore_index: Optional[int] = None,
                 normalize: Optional[Literal["none", "true", "pred", "all"]] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(**kwargs)
        if validate_args:
            F._multiclass_confusion_matrix_arg_validation(num_classes, ignore_index, normalize)
        self.num_classes = num_classes
        self.ignore_index = ignore_index
        self.normalize = normalize
        self.validate_args = validate_args
        self.add_state("confmat", torch.zeros(num_classes, num_classes, dtype=torch.long), dist_reduce_fx="sum")

    def update(self, preds: Tensor, target: Tensor) -> None:
        fv = C.FusedValidation(self.validate_args, preds, target)
        if self.validate_args:
            with fv:
                F._multiclass_confusion_matrix_tensor_validation(preds, target, self.num_classes, self.ignore_index)

        def step(out: Tensor) -> None:
            F._multiclass_confusion_matrix_update_fused(preds, target, self.num_classes, self.ignore_index, out=out)

        step(self.confmat)
        fv.finish(preds.device, self.num_classes, undo=lambda: C.rollback(self.confmat, step), owner=self)

    def compute(self) -> Tensor:
        self._check_device_status(self.num_classes)
        return F._multiclass_confusion_matrix_compute(self.confmat, self.normalize)


class MultilabelConfusionMatrix(_ConfmatBase):
    """(L,2,2) confusion matrices, one per label."""

    def __init__(self, num_labels: int, threshold: float = 0.5, ignore_index: Optional[int] = None,
                 normalize: Optional[Literal["none", "true", "pred", "all"]] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(**kwargs)
        if validate_args:
            F._multilabel_confusion_matrix_arg_validation(num_labels, threshold, ignore_index, normalize)
        self.num_labels = num_labels
        self.threshold = threshold
        self.ignore_index = ignore_index
        self.normalize = normalize
        self.validate_args = validate_args
        self.add_state("confmat", torch.zeros(num_labels, 2, 2, dtype=torch.long), dist_reduce_fx="sum")

    def update(self, preds: Tensor, target: Tensor) -> None:
        fv = C.FusedValidation(self.validate_args, preds, target)
        if self.validate_args:
            with fv:
                F._multilabel_confusion_matrix_tensor_validation(preds, target, self.num_labels, self.ignore_index)

        def step(out: Tensor) -> None:
            F._multilabel_confusion_matrix_update_fused(preds, target, self.num_labels, self.threshold,
                                                        self.ignore_index, out=out)

        step(self.confmat)
        fv.finish(preds.device, undo=lambda: C.rollback(self.confmat, step), owner=self)

    def compute(self) -> Tensor:
        return F._multilabel_confusion_matrix_compute(self.confmat, self.normalize)


class ConfusionMatrix(_ClassificationTaskWrapper):
    """`ConfusionMatrix(task=...)` -> Binary / Multiclass / Multilabel variant."""

    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], threshold: float = 0.5,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        task = ClassificationTask.from_str(task)
        kwargs.update({"normalize": normalize, "ignore_index": ignore_index, "validate_args": validate_args})
        if task == ClassificationTask.BINARY:
            return BinaryConfusionMatrix(threshold, **kwargs)
        if task == ClassificationTask.MULTICLASS:
            if not isinstance(num_classes, int):
                raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
            return MulticlassConfusionMatrix(num_classes, **kwargs)
        if task == ClassificationTask.MULTILABEL:
            if not isinstance(num_labels, int):
                raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
            return MultilabelConfusionMatrix(num_labels, threshold, **kwargs)
        raise ValueError(f"Task {task} not supported!")
