"""Matthews correlation coefficient from the confusion-matrix state
(reference: functional/classification/matthews_corrcoef.py:37-78 reducer; the update is `_*_confusion_matrix_update`,
K1 / K2)."""
from importlib import import_module
from typing import Optional

import torch
from torch import Tensor
from typing_extensions import Literal

CM = import_module("torchmetrics_b200.functional.classification.confusion_matrix")
C = import_module("torchmetrics_b200.functional.classification._checks")
from torchmetrics_b200.utilities.enums import ClassificationTask


def _matthews_corrcoef_reduce(confmat: Tensor) -> Tensor:
    """(C,C) / (2,2) / (L,2,2) un-normalised confusion matrix -> MCC, with the reference's degenerate-case rules."""
    cm = confmat.sum(0) if confmat.ndim == 3 else confmat  # multilabel: pool the labels into one 2x2 table
    binary = cm.numel() == 4
    if binary:
        tn, fp, fn, tp = cm.reshape(-1)
        if tp + tn != 0 and fp + fn == 0:
            return torch.tensor(1.0, dtype=cm.dtype, device=cm.device)
        if tp + tn == 0 and fp + fn != 0:
            return torch.tensor(-1.0, dtype=cm.dtype, device=cm.device)
    tk = cm.sum(dim=-1).float()  # per true class
    pk = cm.sum(dim=-2).float()  # per predicted class
    c = torch.trace(cm).float()
    s = cm.sum().float()
    cov_tp = c * s - sum(tk * pk)
    cov_pp = s ** 2 - sum(pk * pk)
    cov_tt = s ** 2 - sum(tk * tk)
    numerator, denom = cov_tp, cov_pp * cov_tt
    if denom == 0 and binary:
        if tp == 0 or tn == 0:
            a = tp + tn
        if fp == 0 or fn == 0:
            b = fp + fn
        eps = torch.tensor(torch.finfo(torch.float32).eps, dtype=torch.float32, device=cm.device)
        numerator = torch.sqrt(eps) * (a - b)
        denom = (tp + fp + eps) * (tp + fn + eps) * (tn + fp + eps) * (tn + fn + eps)
    elif denom == 0:
        return torch.tensor(0, dtype=cm.dtype, device=cm.device)
    return numerator / torch.sqrt(denom)


def binary_matthews_corrcoef(preds: Tensor, target: Tensor, threshold: float = 0.5,
                             ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        CM._binary_confusion_matrix_arg_validation(threshold, ignore_index)
        with fv:
            CM._binary_confusion_matrix_tensor_validation(preds, target, ignore_index)
    confmat = CM._binary_confusion_matrix_update_fused(preds, target, threshold, ignore_index)
    fv.finish(preds.device, 2, errors=True)
    return _matthews_corrcoef_reduce(confmat)


def multiclass_matthews_corrcoef(preds: Tensor, target: Tensor, num_classes: int, ignore_index: Optional[int] = None,
                                 validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        CM._multiclass_confusion_matrix_arg_validation(num_classes, ignore_index)
        with fv:
            CM._multiclass_confusion_matrix_tensor_validation(preds, target, num_classes, ignore_index)
    confmat = CM._multiclass_confusion_matrix_update_fused(preds, target, num_classes, ignore_index)
    fv.finish(preds.device, num_classes, errors=True)
    return _matthews_corrcoef_reduce(confmat)


def multilabel_matthews_corrcoef(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                                 ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        CM._multilabel_confusion_matrix_arg_validation(num_labels, threshold, ignore_index)
        with fv:
            CM._multilabel_confusion_matrix_tensor_validation(preds, target, num_labels, ignore_index)
    confmat = CM._multilabel_confusion_matrix_update_fused(preds, target, num_labels, threshold, ignore_index)
    fv.finish(preds.device)
    return _matthews_corrcoef_reduce(confmat)


def matthews_corrcoef(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
                      threshold: float = 0.5, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                      ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_matthews_corrcoef(preds, target, threshold, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_matthews_corrcoef(preds, target, num_classes, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_matthews_corrcoef(preds, target, num_labels, threshold, ignore_index, validate_args)
    raise ValueError(f"Not handled value: {task}")
