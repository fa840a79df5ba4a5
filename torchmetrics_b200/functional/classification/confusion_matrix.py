"""Confusion matrices (reference: functional/classification/confusion_matrix.py).

Seams kept with the reference's names and tensor-in / tensor-out contracts: `_*_arg_validation`,
`_*_tensor_validation`, `_*_format`, `_*_update`, `_*_compute`.  The `_update` seams run the K1 / K2 CUDA kernels
(include/tm_b200.h).  `_*_update_fused` additionally folds the per-element work of `_format` (ignore_index filter,
all-or-nothing sigmoid, thresholding) into the same kernel pass and can accumulate straight into a metric state;
the public functions and the Metric classes use those.
"""
from importlib import import_module
from typing import Optional, Tuple

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200 import _native
C = import_module("torchmetrics_b200.functional.classification._checks")
from torchmetrics_b200.utilities.checks import _check_same_shape
from torchmetrics_b200.utilities.enums import ClassificationTask
from torchmetrics_b200.utilities.prints import rank_zero_warn

_NORMALIZE = ("true", "pred", "all", "none", None)


def _confusion_matrix_reduce(confmat: Tensor, normalize: Optional[Literal["true", "pred", "all", "none"]] = None) -> Tensor:
    """Optional row / column / global normalisation (confusion_matrix.py:26-59); NaN rows become 0 with a warning."""
    if normalize not in _NORMALIZE:
        raise ValueError(f"Argument `normalize` needs to one of the following: {_NORMALIZE}")
    if normalize is None or normalize == "none":
        return confmat
    confmat = confmat if confmat.is_floating_point() else confmat.float()
    dims = {"true": -1, "pred": -2, "all": [-2, -1]}[normalize]
    confmat = confmat / confmat.sum(dim=dims, keepdim=True)
    nan_mask = torch.isnan(confmat)
    n_nan = int(nan_mask.sum())
    if n_nan:
        confmat = torch.where(nan_mask, torch.zeros_like(confmat), confmat)
        rank_zero_warn(f"{n_nan} NaN values found in confusion matrix have been replaced with zeros.")
    return confmat


# ======================================================================================= binary
def _binary_confusion_matrix_arg_validation(threshold: float = 0.5, ignore_index: Optional[int] = None,
                                            normalize: Optional[str] = None) -> None:
    C.check_threshold(threshold)
    C.check_ignore_index(ignore_index)
    if normalize not in _NORMALIZE:
        raise ValueError(f"Expected argument `normalize` to be one of {_NORMALIZE}, but got {normalize}.")


def _binary_confusion_matrix_tensor_validation(preds: Tensor, target: Tensor, ignore_index: Optional[int] = None) -> None:
    _check_same_shape(preds, target)
    C.check_binary_values(target, "target", ignore_index)
    if not preds.is_floating_point():
        C.check_binary_values(preds, "preds")


def _binary_confusion_matrix_format(preds: Tensor, target: Tensor, threshold: float = 0.5,
                                    ignore_index: Optional[int] = None, convert_to_labels: bool = True
                                    ) -> Tuple[Tensor, Tensor]:
    """Reference-contract formatting on the device (confusion_matrix.py:118-146)."""
    preds, target = preds.flatten(), target.flatten()
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    if preds.is_floating_point():
        if not C.in_unit_interval(preds):
            preds = preds.sigmoid()
        if convert_to_labels:
            preds = preds > threshold
    return preds, target


def _binary_confusion_matrix_update(preds: Tensor, target: Tensor) -> Tensor:
    """(2,2) int64 counts, [target][pred] (confusion_matrix.py:149-153) — K1 with C = 2."""
    cm = torch.zeros(2, 2, dtype=torch.int64, device=preds.device)
    _native.confmat_update_(cm, preds, target, 2)
    return cm


def _binary_confusion_matrix_update_fused(preds: Tensor, target: Tensor, threshold: float = 0.5,
                                          ignore_index: Optional[int] = None, out: Optional[Tensor] = None) -> Tensor:
    """format + update in one pass.  Float scores: K2 (tp/fp/tn/fn with fused sigmoid rule and threshold)
    rearranged into [[tn, fp], [fn, tp]]; label preds: K1 with the ignore filter in the kernel."""
    dev = preds.device
    if preds.is_floating_point():
        stats = torch.zeros(4, 1, dtype=torch.int64, device=dev)
        # this _format drops ignored samples before the [0,1] range check (:133-142)
        _native.stat_scores_update_(stats, preds, target, 1, threshold, ignore_index, _native.PROB_AUTO,
                                    range_check_ignored=False)
        tp, fp, tn, fn = stats[:, 0]
        cm = torch.stack([torch.stack([tn, fp]), torch.stack([fn, tp])])
        if out is None:
            return cm
        out += cm
        return out
    cm = out if out is not None else torch.zeros(2, 2, dtype=torch.int64, device=dev)
    _native.confmat_update_(cm, preds, target, 2, ignore_index)
    return cm


def _binary_confusion_matrix_compute(confmat: Tensor, normalize: Optional[str] = None) -> Tensor:
    return _confusion_matrix_reduce(confmat, normalize)


def binary_confusion_matrix(preds: Tensor, target: Tensor, threshold: float = 0.5,
                            normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
                            ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        _binary_confusion_matrix_arg_validation(threshold, ignore_index, normalize)
        with fv:
            _binary_confusion_matrix_tensor_validation(preds, target, ignore_index)
    confmat = _binary_confusion_matrix_update_fused(preds, target, threshold, ignore_index)
    fv.finish(preds.device, 2, errors=True)
    return _binary_confusion_matrix_compute(confmat, normalize)


# ======================================================================================= multiclass
def _multiclass_confusion_matrix_arg_validation(num_classes: int, ignore_index: Optional[int] = None,
                                                normalize: Optional[str] = None) -> None:
    C.check_num("num_classes", num_classes)
    C.check_ignore_index(ignore_index)
    if normalize not in _NORMALIZE:
        raise ValueError(f"Expected argument `normalize` to be one of {_NORMALIZE}, but got {normalize}.")


def _multiclass_confusion_matrix_tensor_validation(preds: Tensor, target: Tensor, num_classes: int,
                                                   ignore_index: Optional[int] = None) -> None:
    C.check_multiclass_shapes(preds, target, num_classes)
    C.check_label_count(target, "target", num_classes, ignore_index)
    if not preds.is_floating_point():
        C.check_label_count(preds, "preds", num_classes)


def _multiclass_confusion_matrix_format(preds: Tensor, target: Tensor, ignore_index: Optional[int] = None,
                                        convert_to_labels: bool = True) -> Tuple[Tensor, Tensor]:
    """Reference-contract formatting (confusion_matrix.py:306-330): argmax, flatten, drop ignored samples."""
    if preds.ndim == target.ndim + 1 and convert_to_labels:
        preds = preds.argmax(dim=1)
    preds = preds.flatten() if convert_to_labels else torch.movedim(preds, 1, -1).reshape(-1, preds.shape[1])
    target = target.flatten()
    if ignore_index is not None:
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    return preds, target


def _multiclass_confusion_matrix_update(preds: Tensor, target: Tensor, num_classes: int) -> Tensor:
    """(C,C) int64 counts, rows = target (confusion_matrix.py:333-337) — K1, no key tensor materialised."""
    cm = torch.zeros(num_classes, num_classes, dtype=torch.int64, device=preds.device)
    _native.confmat_update_(cm, preds, target, num_classes)
    return cm


def _multiclass_confusion_matrix_update_fused(preds: Tensor, target: Tensor, num_classes: int,
                                              ignore_index: Optional[int] = None, out: Optional[Tensor] = None) -> Tensor:
    """argmax (when preds are scores) + ignore filter + histogram; accumulates into `out` when given."""
    if preds.ndim == target.ndim + 1:
        if preds.ndim == 2 and preds.is_cuda and preds.is_floating_point():
            preds = _native.rows_argmax(preds)  # int32 ids, same first-maximum / NaN rule as torch.argmax
        else:
            preds = preds.argmax(dim=1)
    cm = out if out is not None else torch.zeros(num_classes, num_classes, dtype=torch.int64, device=preds.device)
    _native.confmat_update_(cm, preds, target, num_classes, ignore_index)
    return cm


def _multiclass_confusion_matrix_compute(confmat: Tensor, normalize: Optional[str] = None) -> Tensor:
    return _confusion_matrix_reduce(confmat, normalize)


def multiclass_confusion_matrix(preds: Tensor, target: Tensor, num_classes: int,
                                normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
                                ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        _multiclass_confusion_matrix_arg_validation(num_classes, ignore_index, normalize)
        with fv:
            _multiclass_confusion_matrix_tensor_validation(preds, target, num_classes, ignore_index)
    confmat = _multiclass_confusion_matrix_update_fused(preds, target, num_classes, ignore_index)
    fv.finish(preds.device, num_classes, errors=True)
    return _multiclass_confusion_matrix_compute(confmat, normalize)


# ======================================================================================= multilabel
def _multilabel_confusion_matrix_arg_validation(num_labels: int, threshold: float = 0.5,
                                                ignore_index: Optional[int] = None,
                                                normalize: Optional[str] = None) -> None:
    C.check_num("num_labels", num_labels)
    C.check_threshold(threshold)
    C.check_ignore_index(ignore_index)
    if normalize not in _NORMALIZE:
        raise ValueError(f"Expected argument `normalize` to be one of {_NORMALIZE}, but got {normalize}.")


def _multilabel_confusion_matrix_tensor_validation(preds: Tensor, target: Tensor, num_labels: int,
                                                   ignore_index: Optional[int] = None) -> None:
    C.check_multilabel_shapes(preds, target, num_labels)
    C.check_binary_values(target, "target", ignore_index)
    if not preds.is_floating_point():
        C.check_binary_values(preds, "preds")


def _multilabel_confusion_matrix_format(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                                        ignore_index: Optional[int] = None, should_threshold: bool = True
                                        ) -> Tuple[Tensor, Tensor]:
    """Reference-contract formatting (confusion_matrix.py:486-518): ignored entries become -4*L."""
    if preds.is_floating_point():
        if not C.in_unit_interval(preds):
            preds = preds.sigmoid()
        if should_threshold:
            preds = preds > threshold
    preds = torch.movedim(preds, 1, -1).reshape(-1, num_labels)
    target = torch.movedim(target, 1, -1).reshape(-1, num_labels)
    if ignore_index is not None:
        preds, target = preds.clone(), target.clone()
        ign = target == ignore_index
        preds[ign] = -4 * num_labels
        target[ign] = -4 * num_labels
    return preds, target


def _label_major(t: Tensor, num_labels: int) -> Tensor:
    """(N, L, ...) -> (N*prod(...), L) row-major, the layout K2 / K3 stream over."""
    if t.ndim == 2:
        return t
    return torch.movedim(t, 1, -1).reshape(-1, num_labels)


def _multilabel_confusion_matrix_update(preds: Tensor, target: Tensor, num_labels: int) -> Tensor:
    """(L,2,2) int64 (confusion_matrix.py:521-526).  Inputs as produced by `_format`: label tensors with ignored
    entries negative — K2 counts an element only when target is 0 or 1."""
    stats = torch.zeros(4, num_labels, dtype=torch.int64, device=preds.device)
    _native.stat_scores_update_(stats, preds.to(torch.int64) if preds.dtype == torch.bool else preds, target,
                                num_labels, 0.5, None, _native.PROB_AS_IS, report_values=False)
    return _stats_to_confmat(stats)


def _stats_to_confmat(stats: Tensor) -> Tensor:
    tp, fp, tn, fn = stats
    return torch.stack([tn, fp, fn, tp], dim=-1).reshape(-1, 2, 2)


def _multilabel_confusion_matrix_update_fused(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                                              ignore_index: Optional[int] = None, out: Optional[Tensor] = None) -> Tensor:
    stats = torch.zeros(4, num_labels, dtype=torch.int64, device=preds.device)
    _native.stat_scores_update_(stats, _label_major(preds, num_labels), _label_major(target, num_labels), num_labels,
                                threshold, ignore_index, _native.PROB_AUTO)
    cm = _stats_to_confmat(stats)
    if out is None:
        return cm
    out += cm
    return out


def _multilabel_confusion_matrix_compute(confmat: Tensor, normalize: Optional[str] = None) -> Tensor:
    return _confusion_matrix_reduce(confmat, normalize)


def multilabel_confusion_matrix(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                                normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
                                ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target)
    if validate_args:
        _multilabel_confusion_matrix_arg_validation(num_labels, threshold, ignore_index, normalize)
        with fv:
            _multilabel_confusion_matrix_tensor_validation(preds, target, num_labels, ignore_index)
    confmat = _multilabel_confusion_matrix_update_fused(preds, target, num_labels, threshold, ignore_index)
    fv.finish(preds.device)
    return _multilabel_confusion_matrix_compute(confmat, normalize)


# ======================================================================================= task wrapper
def confusion_matrix(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
                     threshold: float = 0.5, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                     normalize: Optional[Literal["true", "pred", "all", "none"]] = None,
                     ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    task = ClassificationTask.from_str(task)
    if task == ClassificationTask.BINARY:
        return binary_confusion_matrix(preds, target, threshold, normalize, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_confusion_matrix(preds, target, num_classes, normalize, ignore_index, validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_confusion_matrix(preds, target, num_labels, threshold, normalize, ignore_index, validate_args)
    raise ValueError(f"Task {task} not supported.")
