"""Shared driver for metrics that are ratios of tp / fp / tn / fn (accuracy, precision, recall, F-beta, ...).

The reference spells the same validate -> format -> update -> reduce sequence out once per metric and task
(accuracy.py:89-436, precision_recall.py:60-742, f_beta.py:60-774); here the sequence lives in three helpers and
each public function only supplies its reducer.
"""
from importlib import import_module
from typing import Callable, Optional

from torch import Tensor

from torchmetrics_b200 import _native
S = import_module("torchmetrics_b200.functional.classification.stat_scores")
C = import_module("torchmetrics_b200.functional.classification._checks")

Reducer = Callable[..., Tensor]  # reducer(tp, fp, tn, fn, average=..., multidim_average=..., multilabel=...)


def binary_from_stats(reduce: Reducer, preds: Tensor, target: Tensor, threshold: float, multidim_average: str,
                      ignore_index: Optional[int], validate_args: bool) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target, multidim_average == "global")
    if validate_args:
        S._binary_stat_scores_arg_validation(threshold, multidim_average, ignore_index)
        with fv:
            S._binary_stat_scores_tensor_validation(preds, target, multidim_average, ignore_index)
    tp, fp, tn, fn = S._binary_stat_scores_update_fused(preds, target, threshold, multidim_average, ignore_index)
    fv.finish(preds.device)
    return reduce(tp, fp, tn, fn, average="binary", multidim_average=multidim_average)


def multiclass_from_stats(reduce: Reducer, preds: Tensor, target: Tensor, num_classes: int, average: Optional[str],
                          top_k: int, multidim_average: str, ignore_index: Optional[int], validate_args: bool) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target,
                           S._multiclass_stat_scores_reports_values(top_k, average, multidim_average))
    if validate_args:
        S._multiclass_stat_scores_arg_validation(num_classes, top_k, average, multidim_average, ignore_index)
        with fv:
            S._multiclass_stat_scores_tensor_validation(preds, target, num_classes, multidim_average, ignore_index)
    tp, fp, tn, fn = S._multiclass_stat_scores_update_fused(preds, target, num_classes, top_k, average,
                                                            multidim_average, ignore_index)
    fv.finish(preds.device, num_classes, errors=True)
    return reduce(tp, fp, tn, fn, average=average, multidim_average=multidim_average)


def multilabel_from_stats(reduce: Reducer, preds: Tensor, target: Tensor, num_labels: int, threshold: float,
                          average: Optional[str], multidim_average: str, ignore_index: Optional[int],
                          validate_args: bool) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target, multidim_average == "global")
    if validate_args:
        S._multilabel_stat_scores_arg_validation(num_labels, threshold, average, multidim_average, ignore_index)
        with fv:
            S._multilabel_stat_scores_tensor_validation(preds, target, num_labels, multidim_average, ignore_index)
    tp, fp, tn, fn = S._multilabel_stat_scores_update_fused(preds, target, num_labels, threshold, multidim_average,
                                                            ignore_index)
    fv.finish(preds.device)
    return reduce(tp, fp, tn, fn, average=average, multidim_average=multidim_average, multilabel=True)


def check_task_int(name: str, value) -> None:
    if not isinstance(value, int):
        raise ValueError(f"`{name}` is expected to be `int` but `{type(value)} was passed.`")
