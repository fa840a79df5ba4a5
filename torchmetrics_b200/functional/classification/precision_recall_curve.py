"""Precision-recall curves, binned and exact (reference: functional/classification/precision_recall_curve.py).

Binned mode (`thresholds` given): the `(T[,C],2,2)` multi-threshold confusion matrix is produced by the K3 kernel in
one pass over the scores — the reference needs 3*T reduction passes (`_update_loop`, :228-250) or an N x T key tensor
(`_update_vectorized`, :210-225).  The kernel compares against the caller's own float32 threshold values.
Exact mode (`thresholds=None`): the state is the raw (preds, target); `_binary_clf_curve` (:28-80) is the sort +
scan step behind every exact curve.
"""
from importlib import import_module
from typing import List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor
from torch.nn import functional as F
from typing_extensions import Literal

from torchmetrics_b200 import _native
C = import_module("torchmetrics_b200.functional.classification._checks")
from torchmetrics_b200.utilities.checks import _check_same_shape
from torchmetrics_b200.utilities.compute import _safe_divide, interp
from torchmetrics_b200.utilities.data import _cumsum
from torchmetrics_b200.utilities.enums import ClassificationTask


# ======================================================================================= exact curve core
def _binary_clf_curve(preds: Tensor, target: Tensor, sample_weights: Optional[Union[Sequence, Tensor]] = None,
                      pos_label: int = 1) -> Tuple[Tensor, Tensor, Tensor]:
    """fps, tps and the distinct score values in descending order (precision_recall_curve.py:28-80).

    Without sample weights the counts are accumulated as exact integers on the device and cast to float32 at the
    end (the reference accumulates in float32 and so rounds above 2**24 positives; SURVEY.md hard part 6).
    """
    with torch.no_grad():
        if preds.ndim > target.ndim:
            preds = preds[:, 0]
        if sample_weights is not None:
            if not isinstance(sample_weights, Tensor):
                sample_weights = torch.tensor(sample_weights, device=preds.device, dtype=torch.float)
            order = torch.argsort(preds, descending=True)
            preds, target, weight = preds[order], target[order], sample_weights[order]
            distinct = torch.where(preds[1:] - preds[:-1])[0]
            idx = F.pad(distinct, [0, 1], value=target.size(0) - 1)
            pos = (target == pos_label).to(torch.long)
            tps = _cumsum(pos * weight, dim=0)[idx]
            fps = _cumsum((1 - pos) * weight, dim=0)[idx]
            return fps, tps, preds[idx]
        from torchmetrics_b200.functional.classification._exact import exact_clf_curve

        fps, tps, thr = exact_clf_curve(preds, target, pos_label)
        return fps.to(torch.float32), tps.to(torch.float32), thr


def _adjust_threshold_arg(thresholds: Optional[Union[int, List[float], Tensor]] = None,
                          device: Optional[torch.device] = None) -> Optional[Tensor]:
    """int -> linspace(0, 1, n), list -> tensor (precision_recall_curve.py:83-91)."""
    if isinstance(thresholds, int):
        return torch.linspace(0, 1, thresholds, device=device)
    if isinstance(thresholds, list):
        return torch.tensor(thresholds, device=device)
    return thresholds


def _kernel_thresholds(thresholds: Tensor) -> Tuple[Tensor, Optional[Tensor]]:
    """Ascending float32 thresholds for K3 plus the permutation that restores the caller's order (None when the
    caller's order is already ascending by construction)."""
    thr = thresholds.detach().to(torch.float32)
    sthr, order = torch.sort(thr, stable=True)
    return sthr.contiguous(), order


def _binned_state(preds: Tensor, target: Tensor, thresholds: Tensor, num_columns: int, class_index: bool,
                  ignore_index: Optional[int], prob_mode: int, out: Optional[Tensor], assume_sorted: bool) -> Tensor:
    T = thresholds.numel()
    shape = (T, 2, 2) if num_columns == 1 and not class_index else (T, num_columns, 2, 2)
    if assume_sorted:
        state = out if out is not None else torch.zeros(shape, dtype=torch.int64, device=preds.device)
        _native.binned_curve_update_(state, preds, target, thresholds.to(torch.float32), num_columns, class_index,
                                     ignore_index, prob_mode)
        return state
    sthr, order = _kernel_thresholds(thresholds)
    tmp = torch.zeros(shape, dtype=torch.int64, device=preds.device)
    _native.binned_curve_update_(tmp, preds, target, sthr, num_columns, class_index, ignore_index, prob_mode)
    fresh = torch.empty_like(tmp)
    fresh[order] = tmp
    if out is None:
        return fresh
    out += fresh
    return out


# ======================================================================================= binary
def _binary_precision_recall_curve_arg_validation(thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                                  ignore_index: Optional[int] = None) -> None:
    C.check_thresholds_arg(thresholds)
    C.check_ignore_index(ignore_index)


def _binary_precision_recall_curve_tensor_validation(preds: Tensor, target: Tensor,
                                                     ignore_index: Optional[int] = None) -> None:
    _check_same_shape(preds, target)
    C.check_curve_target_dtype(target)
    C.check_float_preds(preds)
    C.check_binary_values(target, "target", ignore_index)


def _binary_precision_recall_curve_format(preds: Tensor, target: Tensor,
                                          thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                          ignore_index: Optional[int] = None
                                          ) -> Tuple[Tensor, Tensor, Optional[Tensor]]:
    """Reference-contract formatting (precision_recall_curve.py:162-187)."""
    preds, target = preds.flatten(), target.flatten()
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    if not C.in_unit_interval(preds):
        preds = preds.sigmoid()
    return preds, target, _adjust_threshold_arg(thresholds, preds.device)


def _binary_precision_recall_curve_update(preds: Tensor, target: Tensor, thresholds: Optional[Tensor]
                                          ) -> Union[Tensor, Tuple[Tensor, Tensor]]:
    """`(T,2,2)` int64 state, or the inputs themselves in exact mode (precision_recall_curve.py:190-207).
    One K3 pass whatever the sample count: no 50 000-sample switch between two algorithms."""
    if thresholds is None:
        return preds, target
    return _binned_state(preds, target, thresholds, 1, False, None, _native.PROB_AS_IS, None, assume_sorted=False)


# kept for callers that name the two reference variants explicitly; both are the same kernel here
_binary_precision_recall_curve_update_vectorized = _binary_precision_recall_curve_update
_binary_precision_recall_curve_update_loop = _binary_precision_recall_curve_update


def _binary_precision_recall_curve_update_fused(preds: Tensor, target: Tensor, thresholds: Tensor,
                                                ignore_index: Optional[int] = None, out: Optional[Tensor] = None,
                                                assume_sorted: bool = False) -> Tensor:
    """_format (flatten, ignore filter, all-or-nothing sigmoid) + _update in one K3 launch pair."""
    return _binned_state(preds.reshape(-1), target.reshape(-1), thresholds, 1, False, ignore_index,
                         _native.PROB_AUTO, out, assume_sorted)


def _binary_precision_recall_curve_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], thresholds: Optional[Tensor],
                                           pos_label: int = 1) -> Tuple[Tensor, Tensor, Tensor]:
    """precision_recall_curve.py:253-283."""
    if isinstance(state, Tensor) and thresholds is not None:
        tps, fps, fns = state[:, 1, 1], state[:, 0, 1], state[:, 1, 0]
        precision = _safe_divide(tps, tps + fps)
        recall = _safe_divide(tps, tps + fns)
        precision = torch.cat([precision, torch.ones(1, dtype=precision.dtype, device=precision.device)])
        recall = torch.cat([recall, torch.zeros(1, dtype=recall.dtype, device=recall.device)])
        return precision, recall, thresholds
    fps, tps, thr = _binary_clf_curve(state[0], state[1], pos_label=pos_label)
    precision = tps / (tps + fps)
    recall = tps / tps[-1]
    precision = torch.cat([precision.flip(0), torch.ones(1, dtype=precision.dtype, device=precision.device)])
    recall = torch.cat([recall.flip(0), torch.zeros(1, dtype=recall.dtype, device=recall.device)])
    return precision, recall, thr.flip(0).detach().clone()


def binary_precision_recall_curve(preds: Tensor, target: Tensor,
                                  thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                  ignore_index: Optional[int] = None, validate_args: bool = True
                                  ) -> Tuple[Tensor, Tensor, Tensor]:
    if validate_args:
        _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
    state, thr = _binary_curve_state(preds, target, thresholds, ignore_index, validate=validate_args)
    return _binary_precision_recall_curve_compute(state, thr)


def _binary_curve_state(preds: Tensor, target: Tensor, thresholds, ignore_index: Optional[int],
                        validate: bool = False):
    """State + threshold tensor for the one-shot functionals (binned -> fused K3, exact -> formatted inputs).
    `validate`: run `_tensor_validation` too — its target-value check rides on the K3 pass in binned mode."""
    fv = C.FusedValidation(validate, preds, target, thresholds is not None)
    if validate:
        with fv:
            _binary_precision_recall_curve_tensor_validation(preds, target, ignore_index)
    if thresholds is None:
        p, t, _ = _binary_precision_recall_curve_format(preds, target, None, ignore_index)
        return (p, t), None
    thr = _adjust_threshold_arg(thresholds, preds.device)
    state = _binary_precision_recall_curve_update_fused(preds, target, thr, ignore_index,
                                                        assume_sorted=isinstance(thresholds, int))
    fv.finish(preds.device, errors=True)
    return state, thr


# ======================================================================================= multiclass
def _multiclass_precision_recall_curve_arg_validation(num_classes: int,
                                                      thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                                      ignore_index: Optional[int] = None,
                                                      average: Optional[Literal["micro", "macro"]] = None) -> None:
    C.check_num("num_classes", num_classes)
    if average not in (None, "micro", "macro"):
        raise ValueError(f"Expected argument `average` to be one of None, 'micro' or 'macro', but got {average}")
    _binary_precision_recall_curve_arg_validation(thresholds, ignore_index)


def _multiclass_precision_recall_curve_tensor_validation(preds: Tensor, target: Tensor, num_classes: int,
                                                         ignore_index: Optional[int] = None) -> None:
    if not preds.ndim == target.ndim + 1:
        raise ValueError(
            f"Expected `preds` to have one more dimension than `target` but got {preds.ndim} and {target.ndim}"
        )
    C.check_curve_target_dtype(target)
    if not preds.is_floating_point():
        raise ValueError(f"Expected `preds` to be a float tensor, but got {preds.dtype}")
    if preds.shape[1] != num_classes:
        raise ValueError(
            "Expected `preds.shape[1]` to be equal to the number of classes but"
            f" got {preds.shape[1]} and {num_classes}."
        )
    if preds.shape[0] != target.shape[0] or preds.shape[2:] != target.shape[1:]:
        raise ValueError(
            "Expected the shape of `preds` should be (N, C, ...) and the shape of `target` should be (N, ...)"
            f" but got {preds.shape} and {target.shape}"
        )
    C.check_label_count(target, "target", num_classes, ignore_index)


def _class_major(preds: Tensor, num_classes: int) -> Tensor:
    """(N, C, ...) -> (N*prod(...), C)."""
    if preds.ndim == 2:
        return preds
    return preds.transpose(0, 1).reshape(num_classes, -1).T


def _multiclass_precision_recall_curve_format(preds: Tensor, target: Tensor, num_classes: int,
                                              thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                              ignore_index: Optional[int] = None,
                                              average: Optional[Literal["micro", "macro"]] = None
                                              ) -> Tuple[Tensor, Tensor, Optional[Tensor]]:
    """Reference-contract formatting (precision_recall_curve.py:423-455)."""
    preds = _class_major(preds, num_classes)
    target = target.flatten()
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    if not C.in_unit_interval(preds):
        preds = preds.softmax(1)
    if average == "micro":
        preds = preds.flatten()
        target = F.one_hot(target, num_classes=num_classes).flatten()
    return preds, target, _adjust_threshold_arg(thresholds, preds.device)


def _multiclass_precision_recall_curve_update(preds: Tensor, target: Tensor, num_classes: int,
                                              thresholds: Optional[Tensor],
                                              average: Optional[Literal["micro", "macro"]] = None
                                              ) -> Union[Tensor, Tuple[Tensor, Tensor]]:
    """`(T,C,2,2)` state from formatted (M, C) probabilities and (M,) class ids (precision_recall_curve.py:458-527)."""
    if thresholds is None:
        return preds, target
    if average == "micro":
        return _binary_precision_recall_curve_update(preds, target, thresholds)
    return _binned_state(preds, target, thresholds, num_classes, True, None, _native.PROB_AS_IS, None, False)


_multiclass_precision_recall_curve_update_vectorized = _multiclass_precision_recall_curve_update
_multiclass_precision_recall_curve_update_loop = _multiclass_precision_recall_curve_update


def _multiclass_precision_recall_curve_update_fused(preds: Tensor, target: Tensor, num_classes: int,
                                                    thresholds: Tensor, ignore_index: Optional[int] = None,
                                                    average: Optional[str] = None, out: Optional[Tensor] = None,
                                                    assume_sorted: bool = False) -> Tensor:
    """One K3 pass with the class-index one-hot and the ignore filter in the kernel.  The kernel also range-checks
    the scores; only when some lie outside [0,1] (logits) is the batch soft-maxed and binned again — the same host
    decision the reference takes with `torch.all` on every batch (:447-448)."""
    dev = preds.device
    p = _class_major(preds, num_classes)
    t = target.reshape(-1)
    T = thresholds.numel()
    per_class = torch.zeros(T, num_classes, 2, 2, dtype=torch.int64, device=dev)
    _binned_state(p, t, thresholds, num_classes, True, ignore_index, _native.PROB_CHECK, per_class, assume_sorted)
    status = _native.read_status(dev)
    if status & _native.STATUS_OUT_OF_UNIT_RANGE:
        if ignore_index is not None:
            keep = t != ignore_index
            p, t = p[keep], t[keep]
        p = p.softmax(1)
        per_class.zero_()
        _binned_state(p, t, thresholds, num_classes, True, None, _native.PROB_AS_IS, per_class, assume_sorted)
        status = _native.read_status(dev)
    if status & _native.STATUS_THRESHOLD_ORDER:
        raise RuntimeError("thresholds handed to the binned-curve kernel must be sorted ascending")
    # micro: one-hot flattening makes every (sample, class) pair a binary sample => sum of the per-class matrices
    state = per_class.sum(1) if average == "micro" else per_class
    if out is None:
        return state
    out += state
    return out


def _multiclass_precision_recall_curve_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], num_classes: int,
                                               thresholds: Optional[Tensor],
                                               average: Optional[Literal["micro", "macro"]] = None):
    """precision_recall_curve.py:530-583."""
    if average == "micro":
        return _binary_precision_recall_curve_compute(state, thresholds)
    if isinstance(state, Tensor) and thresholds is not None:
        tps, fps, fns = state[:, :, 1, 1], state[:, :, 0, 1], state[:, :, 1, 0]
        precision = _safe_divide(tps, tps + fps)
        recall = _safe_divide(tps, tps + fns)
        precision = torch.cat([precision, torch.ones(1, num_classes, dtype=precision.dtype, device=precision.device)]).T
        recall = torch.cat([recall, torch.zeros(1, num_classes, dtype=recall.dtype, device=recall.device)]).T
        thres = thresholds
        tensor_state = True
    else:
        precision_list, recall_list, thres_list = [], [], []
        for i in range(num_classes):
            p, r, th = _binary_precision_recall_curve_compute((state[0][:, i], state[1]), None, pos_label=i)
            precision_list.append(p)
            recall_list.append(r)
            thres_list.append(th)
        tensor_state = False
    if average == "macro":
        thres = thres.repeat(num_classes) if tensor_state else torch.cat(thres_list, 0)
        thres = thres.sort().values
        mean_precision = precision.flatten() if tensor_state else torch.cat(precision_list, 0)
        mean_precision = mean_precision.sort().values
        mean_recall = torch.zeros_like(mean_precision)
        for i in range(num_classes):
            mean_recall += interp(mean_precision, precision[i] if tensor_state else precision_list[i],
                                  recall[i] if tensor_state else recall_list[i])
        mean_recall /= num_classes
        return mean_precision, mean_recall, thres
    if tensor_state:
        return precision, recall, thres
    return precision_list, recall_list, thres_list


def _multiclass_curve_state(preds: Tensor, target: Tensor, num_classes: int, thresholds, ignore_index, average=None,
                            validate: bool = False):
    fv = C.FusedValidation(validate, preds, target, thresholds is not None)
    if validate:
        with fv:
            _multiclass_precision_recall_curve_tensor_validation(preds, target, num_classes, ignore_index)
    if thresholds is None:
        p, t, _ = _multiclass_precision_recall_curve_format(preds, target, num_classes, None, ignore_index, average)
        return (p, t), None
    thr = _adjust_threshold_arg(thresholds, preds.device)
    state = _multiclass_precision_recall_curve_update_fused(preds, target, num_classes, thr, ignore_index, average,
                                                            assume_sorted=isinstance(thresholds, int))
    fv.finish(preds.device)
    return state, thr


def multiclass_precision_recall_curve(preds: Tensor, target: Tensor, num_classes: int,
                                      thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                      average: Optional[Literal["micro", "macro"]] = None,
                                      ignore_index: Optional[int] = None, validate_args: bool = True):
    if validate_args:
        _multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index, average)
    state, thr = _multiclass_curve_state(preds, target, num_classes, thresholds, ignore_index, average,
                                         validate=validate_args)
    return _multiclass_precision_recall_curve_compute(state, num_classes, thr, average)


# ======================================================================================= multilabel
def _multilabel_precision_recall_curve_arg_validation(num_labels: int,
                                                      thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                                      ignore_index: Optional[int] = None) -> None:
    _multiclass_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)


def _multilabel_precision_recall_curve_tensor_validation(preds: Tensor, target: Tensor, num_labels: int,
                                                         ignore_index: Optional[int] = None) -> None:
    _binary_precision_recall_curve_tensor_validation(preds, target, ignore_index)
    if preds.shape[1] != num_labels:
        raise ValueError(
            "Expected both `target.shape[1]` and `preds.shape[1]` to be equal to the number of labels"
            f" but got {preds.shape[1]} and expected {num_labels}"
        )


def _multilabel_precision_recall_curve_format(preds: Tensor, target: Tensor, num_labels: int,
                                              thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                              ignore_index: Optional[int] = None
                                              ) -> Tuple[Tensor, Tensor, Optional[Tensor]]:
    """Reference-contract formatting (precision_recall_curve.py:739-766)."""
    preds = preds.transpose(0, 1).reshape(num_labels, -1).T
    target = target.transpose(0, 1).reshape(num_labels, -1).T
    if not C.in_unit_interval(preds):
        preds = preds.sigmoid()
    thresholds = _adjust_threshold_arg(thresholds, preds.device)
    if ignore_index is not None and thresholds is not None:
        preds, target = preds.clone(), target.clone()
        ign = target == ignore_index
        fill = -4 * num_labels * len(thresholds)
        preds[ign] = fill
        target[ign] = fill
    return preds, target, thresholds


def _multilabel_precision_recall_curve_update(preds: Tensor, target: Tensor, num_labels: int,
                                              thresholds: Optional[Tensor]
                                              ) -> Union[Tensor, Tuple[Tensor, Tensor]]:
    """`(T,L,2,2)` state from formatted inputs (precision_recall_curve.py:769-793); entries whose target is not 0/1
    (the reference's negative fill for ignored ones) are skipped via the kernel's ignore filter."""
    if thresholds is None:
        return preds, target
    fill = -4 * num_labels * len(thresholds)
    return _binned_state(preds.contiguous(), target.contiguous(), thresholds, num_labels, False, fill,
                         _native.PROB_AS_IS, None, False)


def _multilabel_precision_recall_curve_update_fused(preds: Tensor, target: Tensor, num_labels: int,
                                                    thresholds: Tensor, ignore_index: Optional[int] = None,
                                                    out: Optional[Tensor] = None, assume_sorted: bool = False) -> Tensor:
    p = preds if preds.ndim == 2 else preds.transpose(0, 1).reshape(num_labels, -1).T
    t = target if target.ndim == 2 else target.transpose(0, 1).reshape(num_labels, -1).T
    return _binned_state(p, t, thresholds, num_labels, False, ignore_index, _native.PROB_AUTO, out, assume_sorted)


def _multilabel_precision_recall_curve_compute(state: Union[Tensor, Tuple[Tensor, Tensor]], num_labels: int,
                                               thresholds: Optional[Tensor], ignore_index: Optional[int] = None):
    """precision_recall_curve.py:796-830."""
    if isinstance(state, Tensor) and thresholds is not None:
        tps, fps, fns = state[:, :, 1, 1], state[:, :, 0, 1], state[:, :, 1, 0]
        precision = _safe_divide(tps, tps + fps)
        recall = _safe_divide(tps, tps + fns)
        precision = torch.cat([precision, torch.ones(1, num_labels, dtype=precision.dtype, device=precision.device)])
        recall = torch.cat([recall, torch.zeros(1, num_labels, dtype=recall.dtype, device=recall.device)])
        return precision.T, recall.T, thresholds
    precision_list, recall_list, thres_list = [], [], []
    for i in range(num_labels):
        p, t = state[0][:, i], state[1][:, i]
        if ignore_index is not None:
            keep = t != ignore_index
            p, t = p[keep], t[keep]
        pr, rc, th = _binary_precision_recall_curve_compute((p, t), None, pos_label=1)
        precision_list.append(pr)
        recall_list.append(rc)
        thres_list.append(th)
    return precision_list, recall_list, thres_list


def _multilabel_curve_state(preds: Tensor, target: Tensor, num_labels: int, thresholds, ignore_index,
                            validate: bool = False):
    fv = C.FusedValidation(validate, preds, target, thresholds is not None)
    if validate:
        with fv:
            _multilabel_precision_recall_curve_tensor_validation(preds, target, num_labels, ignore_index)
    if thresholds is None:
        p, t, _ = _multilabel_precision_recall_curve_format(preds, target, num_labels, None, ignore_index)
        return (p, t), None
    thr = _adjust_threshold_arg(thresholds, preds.device)
    state = _multilabel_precision_recall_curve_update_fused(preds, target, num_labels, thr, ignore_index,
                                                            assume_sorted=isinstance(thresholds, int))
    fv.finish(preds.device, errors=True)
    return state, thr


def multilabel_precision_recall_curve(preds: Tensor, target: Tensor, num_labels: int,
                                      thresholds: Optional[Union[int, List[float], Tensor]] = None,
                                      ignore_index: Optional[int] = None, validate_args: bool = True):
    if validate_args:
        _multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
    state, thr = _multilabel_curve_state(preds, target, num_labels, thresholds, ignore_index, validate=validate_args)
    return _multilabel_precision_recall_curve_compute(state, num_labels, thr, ignore_index)


# ======================================================================================= task wrapper
def precision_recall_curve(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
                           thresholds: Optional[Union[int, List[float], Tensor]] = None,
                           num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                           ignore_index: Optional[int] = None, validate_args: bool = True):
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_precision_recall_curve(preds, target, thresholds, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_precision_recall_curve(preds, target, num_classes, thresholds, None, ignore_index,
                                                 validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_precision_recall_curve(preds, target, num_labels, thresholds, ignore_index, validate_args)
    raise ValueError(f"Task {task} not supported.")
