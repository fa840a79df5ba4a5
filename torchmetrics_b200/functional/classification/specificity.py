"""Specificity = tn / (tn + fp) (reference: functional/classification/specificity.py:37-57 reducer; the update is the
stat-scores one, i.e. K2 / K1c)."""
from importlib import import_module
from typing import Optional

from torch import Tensor
from typing_extensions import Literal

D = import_module("torchmetrics_b200.functional.classification._derived")
from torchmetrics_b200.utilities.compute import _adjust_weights_safe_divide, _fused_stat_reduce, _safe_divide
from torchmetrics_b200.utilities.enums import ClassificationTask


def _specificity_reduce(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str],
                        multidim_average: str = "global", multilabel: bool = False) -> Tensor:
    fused = _fused_stat_reduce("specificity", tp, fp, tn, fn, average, multidim_average, multilabel)
    if fused is not None:
        return fused
    if average == "micro":
        d = 0 if multidim_average == "global" else 1
        tn, fp = tn.sum(dim=d), fp.sum(dim=d)
    score = _safe_divide(tn, tn + fp)
    if average in ("binary", "micro"):
        return score
    return _adjust_weights_safe_divide(score, average, multilabel, tp, fp, fn)


def binary_specificity(preds: Tensor, target: Tensor, threshold: float = 0.5,
                       multidim_average: Literal["global", "samplewise"] = "global",
                       ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    return D.binary_from_stats(_specificity_reduce, preds, target, threshold, multidim_average, ignore_index,
                               validate_args)


def multiclass_specificity(preds: Tensor, target: Tensor, num_classes: int,
                           average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro", top_k: int = 1,
                           multidim_average: Literal["global", "samplewise"] = "global",
                           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    return D.multiclass_from_stats(_specificity_reduce, preds, target, num_classes, average, top_k, multidim_average,
                                   ignore_index, validate_args)


def multilabel_specificity(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                           average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                           multidim_average: Literal["global", "samplewise"] = "global",
                           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    return D.multilabel_from_stats(_specificity_reduce, preds, target, num_labels, threshold, average,
                                   multidim_average, ignore_index, validate_args)


def specificity(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
                threshold: float = 0.5, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
                multidim_average: Optional[Literal["global", "samplewise"]] = "global", top_k: Optional[int] = 1,
                ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_specificity(preds, target, threshold, multidim_average, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        D.check_task_int("num_classes", num_classes)
        D.check_task_int("top_k", top_k)
        return multiclass_specificity(preds, target, num_classes, average, top_k, multidim_average, ignore_index,
                                      validate_args)
    if task == ClassificationTask.MULTILABEL:
        D.check_task_int("num_labels", num_labels)
        return multilabel_specificity(preds, target, num_labels, threshold, average, multidim_average, ignore_index,
                                      validate_args)
    raise ValueError(f"Not handled value: {task}")
