"""Cohen's kappa from the confusion-matrix state
(reference: functional/classification/cohen_kappa.py:33-54 reducer; the update is `_*_confusion_matrix_update`, K1 / K2)."""
from importlib import import_module
from typing import Optional

import torch
from torch import Tensor
from typing_extensions import Literal

CM = import_module("torchmetrics_b200.functional.classification.confusion_matrix")
C = import_module("torchmetrics_b200.functional.classification._checks")
from torchmetrics_b200.utilities.enums import ClassificationTaskNoMultilabel

_WEIGHTS = ("linear", "quadratic", "none", None)


def _cohen_kappa_reduce(confmat: Tensor, weights: Optional[str] = None) -> Tensor:
    """1 - sum(W * observed) / sum(W * expected), expected = outer(row sums, column sums) / total."""
    cm = confmat if confmat.is_floating_point() else confmat.float()
    n = cm.shape[0]
    col = cm.sum(dim=0, keepdim=True)
    row = cm.sum(dim=1, keepdim=True)
    expected = row @ col / col.sum()
    if weights is None or weights == "none":
        w = torch.ones_like(cm)
        w.fill_diagonal_(0)
    elif weights in ("linear", "quadratic"):
        idx = torch.arange(n, dtype=cm.dtype, device=cm.device)
        diff = idx.unsqueeze(0) - idx.unsqueeze(1)
        w = diff.abs() if weights == "linear" else diff.pow(2.0)
    else:
        raise ValueError(
            f"Received {weights} for argument ``weights`` but should be either None, 'linear' or 'quadratic'"
        )
    return 1 - torch.sum(w * cm) / torch.sum(w * expected)


def _check_weights(weights: Optional[str]) -> None:
    if weights not in _WEIGHTS:
        raise ValueError(f"Expected argument `weight` to be one of {_WEIGHTS}, but got {weights}.")


def binary_cohen_kappa(preds: Tensor, target: Tensor, threshold: float = 0.5,
                       weights: Optional[Literal["linear", "quadratic", "none"]] = None,
                       ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        CM._binary_confusion_matrix_arg_validation(threshold, ignore_index)
        _check_weights(weights)
        with fv:
            CM._binary_confusion_matrix_tensor_validation(preds, target, ignore_index)
    confmat = CM._binary_confusion_matrix_update_fused(preds, target, threshold, ignore_index)
    fv.finish(preds.device, 2, errors=True)
    return _cohen_kappa_reduce(confmat, weights)


def multiclass_cohen_kappa(preds: Tensor, target: Tensor, num_classes: int,
                           weights: Optional[Literal["linear", "quadratic", "none"]] = None,
                           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        CM._multiclass_confusion_matrix_arg_validation(num_classes, ignore_index)
        _check_weights(weights)
        with fv:
            CM._multiclass_confusion_matrix_tensor_validation(preds, target, num_classes, ignore_index)
    confmat = CM._multiclass_confusion_matrix_update_fused(preds, target, num_classes, ignore_index)
    fv.finish(preds.device, num_classes, errors=True)
    return _cohen_kappa_reduce(confmat, weights)


def cohen_kappa(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass"], threshold: float = 0.5,
                num_classes: Optional[int] = None, weights: Optional[Literal["linear", "quadratic", "none"]] = None,
                ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    task = ClassificationTaskNoMultilabel.from_str(task)
    if task == ClassificationTaskNoMultilabel.BINARY:
        return binary_cohen_kappa(preds, target, threshold, weights, ignore_index, validate_args)
    if task == ClassificationTaskNoMultilabel.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_cohen_kappa(preds, target, num_classes, weights, ignore_index, validate_args)
    raise ValueError(f"Not handled value: {task}")
