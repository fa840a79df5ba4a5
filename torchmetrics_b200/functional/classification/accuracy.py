"""Accuracy (reference: functional/classification/accuracy.py)."""
from importlib import import_module
from typing import Optional

from torch import Tensor
from typing_extensions import Literal

D = import_module("torchmetrics_b200.functional.classification._derived")
from torchmetrics_b200.utilities.compute import _adjust_weights_safe_divide, _fused_stat_reduce, _safe_divide
from torchmetrics_b200.utilities.enums import ClassificationTask


def _accuracy_reduce(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str],
                     multidim_average: str = "global", multilabel: bool = False) -> Tensor:
    """(tp+tn)/all for binary and multilabel, tp/(tp+fn) per class for multiclass (accuracy.py:37-86)."""
    # (tp + tn) / all for binary and multilabel states, tp / (tp + fn) for multiclass ones
    fused = _fused_stat_reduce("accuracy" if multilabel or average == "binary" else "recall", tp, fp, tn, fn, average,
                               multidim_average, multilabel)
    if fused is not None:
        return fused
    if average == "binary":
        return _safe_divide(tp + tn, tp + tn + fp + fn)
    if average == "micro":
        d = 0 if multidim_average == "global" else 1
        tp, fn = tp.sum(dim=d), fn.sum(dim=d)
        if multilabel:
            fp, tn = fp.sum(dim=d), tn.sum(dim=d)
            return _safe_divide(tp + tn, tp + tn + fp + fn)
        return _safe_divide(tp, tp + fn)
    score = _safe_divide(tp + tn, tp + tn + fp + fn) if multilabel else _safe_divide(tp, tp + fn)
    return _adjust_weights_safe_divide(score, average, multilabel, tp, fp, fn)


def binary_accuracy(preds: Tensor, target: Tensor, threshold: float = 0.5,
                    multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                    validate_args: bool = True) -> Tensor:
    return D.binary_from_stats(_accuracy_reduce, preds, target, threshold, multidim_average, ignore_index, validate_args)


def multiclass_accuracy(preds: Tensor, target: Tensor, num_classes: int,
                        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro", top_k: int = 1,
                        multidim_average: Literal["global", "samplewise"] = "global",
                        ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    return D.multiclass_from_stats(_accuracy_reduce, preds, target, num_classes, average, top_k, multidim_average,
                                   ignore_index, validate_args)


def multilabel_accuracy(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                        multidim_average: Literal["global", "samplewise"] = "global",
                        ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    return D.multilabel_from_stats(_accuracy_reduce, preds, target, num_labels, threshold, average, multidim_average,
                                   ignore_index, validate_args)


def accuracy(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
             threshold: float = 0.5, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
             average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
             multidim_average: Literal["global", "samplewise"] = "global", top_k: Optional[int] = 1,
             ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_accuracy(preds, target, threshold, multidim_average, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        D.check_task_int("num_classes", num_classes)
        D.check_task_int("top_k", top_k)
        return multiclass_accuracy(preds, target, num_classes, average, top_k, multidim_average, ignore_index,
                                   validate_args)
    if task == ClassificationTask.MULTILABEL:
        D.check_task_int("num_labels", num_labels)
        return multilabel_accuracy(preds, target, num_labels, threshold, average, multidim_average, ignore_index,
                                   validate_args)
    raise ValueError(f"Not handled value: {task}")
