"""ROC curves (reference: functional/classification/roc.py)."""
from importlib import import_module
from typing import List, Optional, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

P = import_module("torchmetrics_b200.functional.classification.precision_recall_curve")
from torchmetrics_b200.utilities.compute import _safe_divide, interp
from torchmetrics_b200.utilities.enums import ClassificationTask
from torchmetrics_b200.utilities.prints import rank_zero_warn


def _binary_roc_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], thresholds: Optional[Tensor],
                        pos_label: int = 1) -> Tuple[Tensor, Tensor, Tensor]:
    """fpr, tpr, thresholds in decreasing-threshold order (roc.py:40-80)."""
    if isinstance(state, Tensor) and thresholds is not None:
        tps, fps, fns, tns = state[:, 1, 1], state[:, 0, 1], state[:, 1, 0], state[:, 0, 0]
        tpr = _safe_divide(tps, tps + fns).flip(0)
        fpr = _safe_divide(fps, fps + tns).flip(0)
        return fpr, tpr, thresholds.flip(0)
    fps, tps, thres = P._binary_clf_curve(preds=state[0], target=state[1], pos_label=pos_label)
    # the curve starts at (0, 0)
    tps = torch.cat([torch.zeros(1, dtype=tps.dtype, device=tps.device), tps])
    fps = torch.cat([torch.zeros(1, dtype=fps.dtype, device=fps.device), fps])
    thres = torch.cat([torch.ones(1, dtype=thres.dtype, device=thres.device), thres])
    if fps[-1] <= 0:
        rank_zero_warn(
            "No negative samples in targets, false positive value should be meaningless."
            " Returning zero tensor in false positive score",
            UserWarning,
        )
        fpr = torch.zeros_like(thres)
    else:
        fpr = fps / fps[-1]
    if tps[-1] <= 0:
        rank_zero_warn(
            "No positive samples in targets, true positive value should be meaningless."
            " Returning zero tensor in true positive score",
            UserWarning,
        )
        tpr = torch.zeros_like(thres)
    else:
        tpr = tps / tps[-1]
    return fpr, tpr, thres


def binary_roc(preds: Tensor, target: Tensor, thresholds: Optional[Union[int, List[float], Tensor]] = None,
               ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor, Tensor]:
    if validate_args:
        P._binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
    state, thr = P._binary_curve_state(preds, target, thresholds, ignore_index, validate=validate_args)
    return _binary_roc_compute(state, thr)


def _multiclass_roc_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], num_classes: int, thresholds: Optional[Tensor],
                            average: Optional[Literal["micro", "macro"]] = None):
    """roc.py:162-204."""
    if average == "micro":
        return _binary_roc_compute(state, thresholds, pos_label=1)
    if isinstance(state, Tensor) and thresholds is not None:
        tps, fps, fns, tns = state[:, :, 1, 1], state[:, :, 0, 1], state[:, :, 1, 0], state[:, :, 0, 0]
        tpr = _safe_divide(tps, tps + fns).flip(0).T
        fpr = _safe_divide(fps, fps + tns).flip(0).T
        thres = thresholds.flip(0)
        tensor_state = True
    else:
        fpr_list, tpr_list, thres_list = [], [], []
        for i in range(num_classes):
            f, t, th = _binary_roc_compute((state[0][:, i], state[1]), None, pos_label=i)
            fpr_list.append(f)
            tpr_list.append(t)
            thres_list.append(th)
        tensor_state = False
    if average == "macro":
        thres = thres.repeat(num_classes) if tensor_state else torch.cat(thres_list, dim=0)
        thres = thres.sort(descending=True).values
        mean_fpr = fpr.flatten() if tensor_state else torch.cat(fpr_list, dim=0)
        mean_fpr = mean_fpr.sort().values
        mean_tpr = torch.zeros_like(mean_fpr)
        for i in range(num_classes):
            mean_tpr += interp(mean_fpr, fpr[i] if tensor_state else fpr_list[i],
                               tpr[i] if tensor_state else tpr_list[i])
        mean_tpr /= num_classes
        return mean_fpr, mean_tpr, thres
    if tensor_state:
        return fpr, tpr, thres
    return fpr_list, tpr_list, thres_list


def multiclass_roc(preds: Tensor, target: Tensor, num_classes: int,
                   thresholds: Optional[Union[int, List[float], Tensor]] = None,
                   average: Optional[Literal["micro", "macro"]] = None, ignore_index: Optional[int] = None,
                   validate_args: bool = True):
    if validate_args:
        P._multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index, average)
    state, thr = P._multiclass_curve_state(preds, target, num_classes, thresholds, ignore_index, average,
                                           validate=validate_args)
    return _multiclass_roc_compute(state, num_classes, thr, average)


def _multilabel_roc_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], num_labels: int, thresholds: Optional[Tensor],
                            ignore_index: Optional[int] = None):
    """roc.py:329-356."""
    if isinstance(state, Tensor) and thresholds is not None:
        tps, fps, fns, tns = state[:, :, 1, 1], state[:, :, 0, 1], state[:, :, 1, 0], state[:, :, 0, 0]
        tpr = _safe_divide(tps, tps + fns).flip(0).T
        fpr = _safe_divide(fps, fps + tns).flip(0).T
        return fpr, tpr, thresholds.flip(0)
    fpr, tpr, thres = [], [], []
    for i in range(num_labels):
        p, t = state[0][:, i], state[1][:, i]
        if ignore_index is not None:
            keep = t != ignore_index
            p, t = p[keep], t[keep]
        f, tp_, th = _binary_roc_compute((p, t), None, pos_label=1)
        fpr.append(f)
        tpr.append(tp_)
        thres.append(th)
    return fpr, tpr, thres


def multilabel_roc(preds: Tensor, target: Tensor, num_labels: int,
                   thresholds: Optional[Union[int, List[float], Tensor]] = None, ignore_index: Optional[int] = None,
                   validate_args: bool = True):
    if validate_args:
        P._multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
    state, thr = P._multilabel_curve_state(preds, target, num_labels, thresholds, ignore_index,
                                           validate=validate_args)
    return _multilabel_roc_compute(state, num_labels, thr, ignore_index)


def roc(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
        thresholds: Optional[Union[int, List[float], Tensor]] = None, num_classes: Optional[int] = None,
        num_labels: Optional[int] = None, ignore_index: Optional[int] = None, validate_args: bool = True):
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_roc(preds, target, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_roc(preds, target, num_classes, thresholds, None, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_roc(preds, target, num_labels, thresholds, ignore_index, validate_args)
    raise ValueError(f"Task {task} not supported.")
