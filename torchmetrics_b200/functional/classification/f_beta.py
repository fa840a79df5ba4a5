"""F-beta and F1 (reference: functional/classification/f_beta.py)."""
from importlib import import_module
from functools import partial
from typing import Optional

from torch import Tensor
from typing_extensions import Literal

D = import_module("torchmetrics_b200.functional.classification._derived")
from torchmetrics_b200.utilities.compute import _adjust_weights_safe_divide, _fused_stat_reduce, _safe_divide
from torchmetrics_b200.utilities.enums import ClassificationTask


def _fbeta_reduce(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, beta: float, average: Optional[str],
                  multidim_average: str = "global", multilabel: bool = False) -> Tensor:
    """(1+b^2) tp / ((1+b^2) tp + b^2 fn + fp) (f_beta.py:37-57)."""
    fused = _fused_stat_reduce("fbeta", tp, fp, tn, fn, average, multidim_average, multilabel, beta)
    if fused is not None:
        return fused
    b2 = beta ** 2
    if average == "binary":
        return _safe_divide((1 + b2) * tp, (1 + b2) * tp + b2 * fn + fp)
    if average == "micro":
        d = 0 if multidim_average == "global" else 1
        tp, fn, fp = tp.sum(dim=d), fn.sum(dim=d), fp.sum(dim=d)
        return _safe_divide((1 + b2) * tp, (1 + b2) * tp + b2 * fn + fp)
    score = _safe_divide((1 + b2) * tp, (1 + b2) * tp + b2 * fn + fp)
    return _adjust_weights_safe_divide(score, average, multilabel, tp, fp, fn)


def _check_beta(beta: float) -> None:
    if not (isinstance(beta, float) and beta > 0):
        raise ValueError(f"Expected argument `beta` to be a float larger than 0, but got {beta}.")


def _reducer(beta: float):
    def reduce(tp, fp, tn, fn, average, multidim_average="global", multilabel=False):
        return _fbeta_reduce(tp, fp, tn, fn, beta, average, multidim_average, multilabel)

    return reduce


def binary_fbeta_score(preds: Tensor, target: Tensor, beta: float, threshold: float = 0.5,
                       multidim_average: Literal["global", "samplewise"] = "global",
                       ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    if validate_args:
        _check_beta(beta)
    return D.binary_from_stats(_reducer(beta), preds, target, threshold, multidim_average, ignore_index, validate_args)


def multiclass_fbeta_score(preds: Tensor, target: Tensor, beta: float, num_classes: int,
                           average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro", top_k: int = 1,
                           multidim_average: Literal["global", "samplewise"] = "global",
                           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    if validate_args:
        _check_beta(beta)
    return D.multiclass_from_stats(_reducer(beta), preds, target, num_classes, average, top_k, multidim_average,
                                   ignore_index, validate_args)


def multilabel_fbeta_score(preds: Tensor, target: Tensor, beta: float, num_labels: int, threshold: float = 0.5,
                           average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                           multidim_average: Literal["global", "samplewise"] = "global",
                           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    if validate_args:
        _check_beta(beta)
    return D.multilabel_from_stats(_reducer(beta), preds, target, num_labels, threshold, average, multidim_average,
                                   ignore_index, validate_args)


def binary_f1_score(preds: Tensor, target: Tensor, threshold: float = 0.5,
                    multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                    validate_args: bool = True) -> Tensor:
    return binary_fbeta_score(preds, target, 1.0, threshold, multidim_average, ignore_index, validate_args)


def multiclass_f1_score(preds: Tensor, target: Tensor, num_classes: int,
                        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro", top_k: int = 1,
                        multidim_average: Literal["global", "samplewise"] = "global",
                        ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    return multiclass_fbeta_score(preds, target, 1.0, num_classes, average, top_k, multidim_average, ignore_index,
                                  validate_args)


def multilabel_f1_score(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                        average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                        multidim_average: Literal["global", "samplewise"] = "global",
                        ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    return multilabel_fbeta_score(preds, target, 1.0, num_labels, threshold, average, multidim_average, ignore_index,
                                  validate_args)


def fbeta_score(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"], beta: float = 1.0,
                threshold: float = 0.5, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
                multidim_average: Optional[Literal["global", "samplewise"]] = "global", top_k: Optional[int] = 1,
                ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    t = ClassificationTask.from_str(task)
    assert multidim_average is not None
    if t == ClassificationTask.BINARY:
        return binary_fbeta_score(preds, target, beta, threshold, multidim_average, ignore_index, validate_args)
    if t == ClassificationTask.MULTICLASS:
        D.check_task_int("num_classes", num_classes)
        D.check_task_int("top_k", top_k)
        return multiclass_fbeta_score(preds, target, beta, num_classes, average, top_k, multidim_average,
                                      ignore_index, validate_args)
    if t == ClassificationTask.MULTILABEL:
        D.check_task_int("num_labels", num_labels)
        return multilabel_fbeta_score(preds, target, beta, num_labels, threshold, average, multidim_average,
                                      ignore_index, validate_args)
    raise ValueError(f"Not handled value: {task}")


def f1_score(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
             threshold: float = 0.5, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
             average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
             multidim_average: Optional[Literal["global", "samplewise"]] = "global", top_k: Optional[int] = 1,
             ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    return fbeta_score(preds, target, task, 1.0, threshold, num_classes, num_labels, average, multidim_average, top_k,
                       ignore_index, validate_args)
