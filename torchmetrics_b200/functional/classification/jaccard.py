"""Jaccard index (intersection over union) from the confusion-matrix state
(reference: functional/classification/jaccard.py:38-94 reducer; the update is `_*_confusion_matrix_update`, K1 / K2)."""
from importlib import import_module
from typing import Optional

import torch
from torch import Tensor
from typing_extensions import Literal

CM = import_module("torchmetrics_b200.functional.classification.confusion_matrix")
C = import_module("torchmetrics_b200.functional.classification._checks")
from torchmetrics_b200.utilities.compute import _safe_divide
from torchmetrics_b200.utilities.enums import ClassificationTask

_AVERAGES = ("binary", "micro", "macro", "weighted", "none", None)


def _jaccard_index_reduce(confmat: Tensor, average: Optional[str], ignore_index: Optional[int] = None) -> Tensor:
    """Un-normalised confusion matrix -> score.  (2,2) binary, (C,C) multiclass, (L,2,2) multilabel."""
    if average not in _AVERAGES:
        raise ValueError(f"The `average` has to be one of {list(_AVERAGES)}, got {average}.")
    cm = confmat.float()
    if average == "binary":
        return cm[1, 1] / (cm[0, 1] + cm[1, 0] + cm[1, 1])
    multilabel = cm.ndim == 3
    drop = ignore_index is not None and 0 <= ignore_index < cm.shape[0]
    if multilabel:
        inter = cm[:, 1, 1]
        union = cm[:, 1, 1] + cm[:, 0, 1] + cm[:, 1, 0]
        support = cm[:, 1, 1] + cm[:, 1, 0]
    else:
        inter = torch.diag(cm)
        union = cm.sum(0) + cm.sum(1) - inter
        support = cm.sum(1)
    if average == "micro":
        inter = inter.sum()
        union = union.sum() - (union[ignore_index] if drop else 0.0)
    score = _safe_divide(inter, union)
    if average in (None, "none", "micro"):
        return score
    if average == "weighted":
        weights = support
    else:
        weights = torch.ones_like(score)
        if drop:
            weights[ignore_index] = 0.0
        if not multilabel:
            weights[cm.sum(1) + cm.sum(0) == 0] = 0.0
    return ((weights * score) / weights.sum()).sum()


def _check_average(average: Optional[str]) -> None:
    allowed = ("micro", "macro", "weighted", "none", None)
    if average not in allowed:
        raise ValueError(f"Expected argument `average` to be one of {allowed}, but got {average}.")


def binary_jaccard_index(preds: Tensor, target: Tensor, threshold: float = 0.5, ignore_index: Optional[int] = None,
                         validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        CM._binary_confusion_matrix_arg_validation(threshold, ignore_index)
        with fv:
            CM._binary_confusion_matrix_tensor_validation(preds, target, ignore_index)
    confmat = CM._binary_confusion_matrix_update_fused(preds, target, threshold, ignore_index)
    fv.finish(preds.device, 2, errors=True)
    return _jaccard_index_reduce(confmat, average="binary")


def multiclass_jaccard_index(preds: Tensor, target: Tensor, num_classes: int,
                             average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                             ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        CM._multiclass_confusion_matrix_arg_validation(num_classes, ignore_index)
        _check_average(average)
        with fv:
            CM._multiclass_confusion_matrix_tensor_validation(preds, target, num_classes, ignore_index)
    confmat = CM._multiclass_confusion_matrix_update_fused(preds, target, num_classes, ignore_index)
    fv.finish(preds.device, num_classes, errors=True)
    return _jaccard_index_reduce(confmat, average=average, ignore_index=ignore_index)


def multilabel_jaccard_index(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                             average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                             ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        CM._multilabel_confusion_matrix_arg_validation(num_labels, threshold, ignore_index)
        _check_average(average)
        with fv:
            CM._multilabel_confusion_matrix_tensor_validation(preds, target, num_labels, ignore_index)
    confmat = CM._multilabel_confusion_matrix_update_fused(preds, target, num_labels, threshold, ignore_index)
    fv.finish(preds.device)
    return _jaccard_index_reduce(confmat, average=average, ignore_index=ignore_index)


def jaccard_index(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
                  threshold: float = 0.5, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                  average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                  ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_jaccard_index(preds, target, threshold, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_jaccard_index(preds, target, num_classes, average, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_jaccard_index(preds, target, num_labels, threshold, average, ignore_index, validate_args)
    raise ValueError(f"Not handled value: {task}")
