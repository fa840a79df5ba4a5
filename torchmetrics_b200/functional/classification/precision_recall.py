"""Precision and recall (reference: functional/classification/precision_recall.py)."""
from importlib import import_module
from functools import partial
from typing import Optional

from torch import Tensor
from typing_extensions import Literal

D = import_module("torchmetrics_b200.functional.classification._derived")
from torchmetrics_b200.utilities.compute import _adjust_weights_safe_divide, _fused_stat_reduce, _safe_divide
from torchmetrics_b200.utilities.enums import ClassificationTask


def _precision_recall_reduce(stat: Literal["precision", "recall"], tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor,
                             average: Optional[str], multidim_average: str = "global",
                             multilabel: bool = False) -> Tensor:
    """tp / (tp + fp) or tp / (tp + fn) (precision_recall.py:37-57)."""
    fused = _fused_stat_reduce(stat, tp, fp, tn, fn, average, multidim_average, multilabel)
    if fused is not None:
        return fused
    other = fp if stat == "precision" else fn
    if average == "binary":
        return _safe_divide(tp, tp + other)
    if average == "micro":
        d = 0 if multidim_average == "global" else 1
        tp, other = tp.sum(dim=d), other.sum(dim=d)
        return _safe_divide(tp, tp + other)
    return _adjust_weights_safe_divide(_safe_divide(tp, tp + other), average, multilabel, tp, fp, fn)


_precision = partial(_precision_recall_reduce, "precision")
_recall = partial(_precision_recall_reduce, "recall")


def _make(reduce, kind):
    def binary(preds: Tensor, target: Tensor, threshold: float = 0.5,
               multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
               validate_args: bool = True) -> Tensor:
        return D.binary_from_stats(reduce, preds, target, threshold, multidim_average, ignore_index, validate_args)

    def multiclass(preds: Tensor, target: Tensor, num_classes: int,
                   average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro", top_k: int = 1,
                   multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                   validate_args: bool = True) -> Tensor:
        return D.multiclass_from_stats(reduce, preds, target, num_classes, average, top_k, multidim_average,
                                       ignore_index, validate_args)

    def multilabel(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                   average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                   multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                   validate_args: bool = True) -> Tensor:
        return D.multilabel_from_stats(reduce, preds, target, num_labels, threshold, average, multidim_average,
                                       ignore_index, validate_args)

    def task_fn(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
                threshold: float = 0.5, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
                multidim_average: Optional[Literal["global", "samplewise"]] = "global", top_k: Optional[int] = 1,
                ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
        assert multidim_average is not None
        t = ClassificationTask.from_str(task)
        if t == ClassificationTask.BINARY:
            return binary(preds, target, threshold, multidim_average, ignore_index, validate_args)
        if t == ClassificationTask.MULTICLASS:
            D.check_task_int("num_classes", num_classes)
            D.check_task_int("top_k", top_k)
            return multiclass(preds, target, num_classes, average, top_k, multidim_average, ignore_index, validate_args)
        if t == ClassificationTask.MULTILABEL:
            D.check_task_int("num_labels", num_labels)
            return multilabel(preds, target, num_labels, threshold, average, multidim_average, ignore_index,
                              validate_args)
        raise ValueError(f"Not handled value: {task}")

    for f, prefix in ((binary, "binary_"), (multiclass, "multiclass_"), (multilabel, "multilabel_"), (task_fn, "")):
        f.__name__ = f.__qualname__ = prefix + kind
    return binary, multiclass, multilabel, task_fn


binary_precision, multiclass_precision, multilabel_precision, precision = _make(_precision, "precision")
binary_recall, multiclass_recall, multilabel_recall, recall = _make(_recall, "recall")
