"""tp / fp / tn / fn statistics (reference: functional/classification/stat_scores.py).

`_binary_stat_scores_update`, `_multilabel_stat_scores_update` (global) run the K2 kernel; the multiclass update
runs K1 (confusion matrix, then diag / row / column sums exactly as stat_scores.py:405-418) or the match-count kernel
for `average="micro"` (:394-404).  `_*_update_fused` folds `_format` into the kernel pass.

`multidim_average="samplewise"` returns per-sample rows instead of O(C) counts: the binary / multilabel forms run the
row-wise K2 kernel (`tmb_stat_scores_rows`, one pass over the raw scores), the multiclass form the row-wise K1c kernel
(`tmb_class_stats_rows`, three per-sample histograms instead of (N, X, C) one-hot tensors).  Only samplewise
`top_k > 1` still follows the reference's one-hot formulation with device-side torch ops (:363-393).
"""
from importlib import import_module
from typing import Optional, Tuple

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200 import _native
C = import_module("torchmetrics_b200.functional.classification._checks")
from torchmetrics_b200.functional.classification.confusion_matrix import _label_major
from torchmetrics_b200.utilities.checks import _check_same_shape
from torchmetrics_b200.utilities.data import select_topk
from torchmetrics_b200.utilities.enums import ClassificationTask

_AVG = ("micro", "macro", "weighted", "none", None)
_MDA = ("global", "samplewise")


def _four_masked_sums(preds: Tensor, target: Tensor, dims) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    eq = target == preds
    tp = (eq & (target == 1)).sum(dims)
    fn = (~eq & (target == 1)).sum(dims)
    fp = (~eq & (target == 0)).sum(dims)
    tn = (eq & (target == 0)).sum(dims)
    return tp, fp, tn, fn


# ======================================================================================= binary
def _binary_stat_scores_arg_validation(threshold: float = 0.5, multidim_average: str = "global",
                                       ignore_index: Optional[int] = None) -> None:
    C.check_threshold(threshold)
    C.check_choice("multidim_average", multidim_average, _MDA)
    C.check_ignore_index(ignore_index)


def _binary_stat_scores_tensor_validation(preds: Tensor, target: Tensor, multidim_average: str = "global",
                                          ignore_index: Optional[int] = None) -> None:
    _check_same_shape(preds, target)
    C.check_binary_values(target, "target", ignore_index)
    if not preds.is_floating_point():
        C.check_binary_values(preds, "preds", what="`preds`")
    if multidim_average != "global" and preds.ndim < 2:
        raise ValueError("Expected input to be at least 2D when multidim_average is set to `samplewise`")


def _binary_stat_scores_format(preds: Tensor, target: Tensor, threshold: float = 0.5,
                               ignore_index: Optional[int] = None) -> Tuple[Tensor, Tensor]:
    """Reference-contract formatting (stat_scores.py:90-117): ignored targets become -1."""
    if preds.is_floating_point():
        if not C.in_unit_interval(preds):
            preds = preds.sigmoid()
        preds = preds > threshold
    preds = preds.reshape(preds.shape[0], -1)
    target = target.reshape(target.shape[0], -1)
    if ignore_index is not None:
        target = torch.where(target == ignore_index, torch.full_like(target, -1), target)
    return preds, target


def _binary_stat_scores_update(preds: Tensor, target: Tensor, multidim_average: str = "global"
                               ) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Counts over formatted label tensors (stat_scores.py:120-131); global -> K2 with one label."""
    if multidim_average == "global":
        stats = torch.zeros(4, 1, dtype=torch.int64, device=preds.device)
        _native.stat_scores_update_(stats, preds, target, 1, 0.5, None, _native.PROB_AS_IS, report_values=False)
        return stats[0, 0], stats[1, 0], stats[2, 0], stats[3, 0]
    rows = _native.stat_scores_rows(preds, target, preds.shape[1] if preds.ndim > 1 else 1)
    return rows[0].squeeze(), rows[1].squeeze(), rows[2].squeeze(), rows[3].squeeze()


def _binary_stat_scores_update_fused(preds: Tensor, target: Tensor, threshold: float = 0.5,
                                     multidim_average: str = "global", ignore_index: Optional[int] = None,
                                     out: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """_format + _update in one K2 pass (global); `out` is an optional (4,1) int64 accumulator."""
    if multidim_average != "global":
        # one row per sample: threshold, sigmoid and ignore filter happen in the row-wise kernel; only the reference's
        # batch-wide range decision (stat_scores.py:104-106) is taken here
        n = preds.shape[0]
        sigmoid = preds.is_floating_point() and not C.in_unit_interval(preds)
        rows = _native.stat_scores_rows(preds, target, preds.numel() // n if n else 0, threshold, ignore_index, sigmoid)
        return rows[0].squeeze(), rows[1].squeeze(), rows[2].squeeze(), rows[3].squeeze()
    stats = out if out is not None else torch.zeros(4, 1, dtype=torch.int64, device=preds.device)
    _native.stat_scores_update_(stats, preds, target, 1, threshold, ignore_index, _native.PROB_AUTO)
    return stats[0, 0], stats[1, 0], stats[2, 0], stats[3, 0]


def _binary_stat_scores_compute(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor,
                                multidim_average: str = "global") -> Tensor:
    return torch.stack([tp, fp, tn, fn, tp + fn], dim=0 if multidim_average == "global" else 1).squeeze()


def binary_stat_scores(preds: Tensor, target: Tensor, threshold: float = 0.5,
                       multidim_average: Literal["global", "samplewise"] = "global",
                       ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target, multidim_average == "global")
    if validate_args:
        _binary_stat_scores_arg_validation(threshold, multidim_average, ignore_index)
        with fv:
            _binary_stat_scores_tensor_validation(preds, target, multidim_average, ignore_index)
    tp, fp, tn, fn = _binary_stat_scores_update_fused(preds, target, threshold, multidim_average, ignore_index)
    fv.finish(preds.device)
    return _binary_stat_scores_compute(tp, fp, tn, fn, multidim_average)


# ======================================================================================= multiclass
def _multiclass_stat_scores_arg_validation(num_classes: int, top_k: int = 1, average: Optional[str] = "macro",
                                           multidim_average: str = "global", ignore_index: Optional[int] = None) -> None:
    C.check_num("num_classes", num_classes)
    if not isinstance(top_k, int) and top_k < 1:
        raise ValueError(f"Expected argument `top_k` to be an integer larger than or equal to 1, but got {top_k}")
    if top_k > num_classes:
        raise ValueError(
            f"Expected argument `top_k` to be smaller or equal to `num_classes` but got {top_k} and {num_classes}"
        )
    C.check_choice("average", average, _AVG)
    C.check_choice("multidim_average", multidim_average, _MDA)
    C.check_ignore_index(ignore_index)


def _multiclass_stat_scores_tensor_validation(preds: Tensor, target: Tensor, num_classes: int,
                                              multidim_average: str = "global",
                                              ignore_index: Optional[int] = None) -> None:
    C.check_multiclass_shapes(preds, target, num_classes, multidim_average)
    C.check_label_count(target, "target", num_classes, ignore_index)
    if not preds.is_floating_point():
        C.check_label_count(preds, "preds", num_classes)


def _multiclass_stat_scores_format(preds: Tensor, target: Tensor, top_k: int = 1) -> Tuple[Tensor, Tensor]:
    """argmax when preds carry a class dimension and top_k == 1; flatten extra dims (stat_scores.py:325-341)."""
    if preds.ndim == target.ndim + 1 and top_k == 1:
        preds = preds.argmax(dim=1)
    preds = preds.reshape(*preds.shape[:2], -1) if top_k != 1 else preds.reshape(preds.shape[0], -1)
    target = target.reshape(target.shape[0], -1)
    return preds, target


def _multiclass_stat_scores_update(preds: Tensor, target: Tensor, num_classes: int, top_k: int = 1,
                                   average: Optional[str] = "macro", multidim_average: str = "global",
                                   ignore_index: Optional[int] = None,
                                   out: Optional[Tensor] = None) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """stat_scores.py:344-419.  `out`: optional (4, C) int64 accumulator for the branches that count per class; the
    returned rows are then its rows (the other branches ignore it and return fresh tensors)."""
    dev = preds.device
    if (multidim_average == "global" and top_k > 1 and preds.is_cuda and preds.is_floating_point()
            and preds.ndim == target.ndim + 1):
        # the reference one-hots the top-k scores and the target and takes four masked sums over (N, X, C) tensors
        # (:363-393); the class ids of the k highest scores are all the kernel needs (K1c, top-k form)
        idx = preds.topk(top_k, dim=1).indices  # same call as select_topk: ties resolve identically
        idx = torch.movedim(idx, 1, -1).reshape(-1, top_k)
        stats = out if out is not None else torch.zeros(4, num_classes, dtype=torch.int64, device=dev)
        _native.class_stats_topk_update_(stats, idx, target.reshape(-1), num_classes, ignore_index)
        return stats[0], stats[1], stats[2], stats[3]
    if (multidim_average == "samplewise" and top_k > 1 and preds.is_cuda and preds.is_floating_point()
            and preds.ndim == target.ndim + 1 and target.ndim >= 2):
        # (N, C, ...) scores: the ids of the k highest scores of every element, laid out (N, X, k), go through the same
        # per-sample histograms (K1c, row-wise top-k form) instead of (N, X, C) one-hot tensors
        n = preds.shape[0]
        idx = torch.movedim(preds.topk(top_k, dim=1).indices, 1, -1).reshape(n, -1, top_k)
        rows = _native.class_stats_rows(idx, target.reshape(n, -1), num_classes, ignore_index, top_k=top_k)
        return rows[0], rows[1], rows[2], rows[3]
    if multidim_average == "samplewise" and top_k == 1 and not preds.is_floating_point():
        # (N, X) label tensors: per sample and class counts from three per-sample histograms (K1c, row-wise form)
        rows = _native.class_stats_rows(preds, target, num_classes, ignore_index)
        return rows[0], rows[1], rows[2], rows[3]
    if multidim_average == "samplewise" or top_k != 1:
        return _multiclass_stat_scores_onehot(preds, target, num_classes, top_k, multidim_average, ignore_index)
    if average == "micro":
        cnt = torch.zeros(2, dtype=torch.int64, device=dev)
        _native.match_count_update_(cnt, preds, target, ignore_index)
        n_valid, tp = cnt[0], cnt[1]
        fp = n_valid - tp
        fn = fp.clone()
        tn = num_classes * n_valid - (fp + fn + tp)
        return tp, fp, tn, fn
    # the reference fills a C x C confusion matrix and reduces it (:395-419); the four vectors only need three
    # length-C histograms (K1c), so nothing of size C*C is allocated, zeroed or summed
    stats = out if out is not None else torch.zeros(4, num_classes, dtype=torch.int64, device=dev)
    _native.class_stats_update_(stats, preds, target, num_classes, ignore_index)
    return stats[0], stats[1], stats[2], stats[3]


def _multiclass_stat_scores_onehot(preds: Tensor, target: Tensor, num_classes: int, top_k: int,
                                   multidim_average: str, ignore_index: Optional[int]):
    """One-hot branch of the reference (stat_scores.py:363-393), device-side torch ops."""
    ignore_in = 0 <= ignore_index <= num_classes - 1 if ignore_index is not None else None
    if ignore_index is not None and not ignore_in:
        preds, target = preds.clone(), target.clone()
        idx = target == ignore_index
        target[idx] = num_classes
        idx = idx.unsqueeze(1).repeat(1, num_classes, 1) if preds.ndim > target.ndim else idx
        preds[idx] = num_classes
    depth = num_classes + 1 if ignore_index is not None and not ignore_in else num_classes
    if top_k > 1:
        preds_oh = torch.movedim(select_topk(preds, topk=top_k, dim=1), 1, -1)
    else:
        preds_oh = torch.nn.functional.one_hot(preds.long(), depth)
    target_oh = torch.nn.functional.one_hot(target.long(), depth)
    if ignore_index is not None:
        if ignore_in:
            target_oh[target == ignore_index, :] = -1
        else:
            preds_oh = preds_oh[..., :-1] if top_k == 1 else preds_oh
            target_oh = target_oh[..., :-1]
            target_oh[target == num_classes, :] = -1
    return _four_masked_sums(preds_oh, target_oh, [0, 1] if multidim_average == "global" else [1])


def _multiclass_stat_scores_reports_values(top_k: int, average: Optional[str], multidim_average: str) -> bool:
    """True when `_multiclass_stat_scores_update` goes through K1, whose kernel validates label values as it counts."""
    return multidim_average == "global" and top_k == 1 and average != "micro"


def _multiclass_stat_scores_update_fused(preds: Tensor, target: Tensor, num_classes: int, top_k: int = 1,
                                         average: Optional[str] = "macro", multidim_average: str = "global",
                                         ignore_index: Optional[int] = None, out: Optional[Tensor] = None):
    if out is not None and tuple(out.shape) != (4, num_classes):
        out = None  # micro-averaged states hold one number each
    if (top_k == 1 and multidim_average == "global" and average != "micro" and preds.ndim == 2 and target.ndim == 1
            and preds.is_cuda and preds.is_floating_point() and preds.shape[1] == num_classes):
        # (N, C) scores: the argmax of `_format` rides on the counting kernel, one warp per row (K1c, argmax form)
        stats = out if out is not None else torch.zeros(4, num_classes, dtype=torch.int64, device=preds.device)
        _native.class_stats_argmax_update_(stats, preds, target, num_classes, ignore_index)
        return stats[0], stats[1], stats[2], stats[3]
    preds, target = _multiclass_stat_scores_format(preds, target, top_k)
    return _multiclass_stat_scores_update(preds, target, num_classes, top_k, average, multidim_average, ignore_index,
                                          out)


def _multiclass_stat_scores_compute(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str] = "macro",
                                    multidim_average: str = "global") -> Tensor:
    """Stack [tp, fp, tn, fn, support] and average over classes (stat_scores.py:422-448)."""
    res = torch.stack([tp, fp, tn, fn, tp + fn], dim=-1)
    sum_dim = 0 if multidim_average == "global" else 1
    if average == "micro":
        return res.sum(sum_dim) if res.ndim > 1 else res
    if average == "macro":
        return res.float().mean(sum_dim)
    if average == "weighted":
        weight = tp + fn
        if multidim_average == "global":
            return (res * (weight / weight.sum()).reshape(*weight.shape, 1)).sum(sum_dim)
        return (res * (weight / weight.sum(-1, keepdim=True)).reshape(*weight.shape, 1)).sum(sum_dim)
    if average is None or average == "none":
        return res
    return None


def multiclass_stat_scores(preds: Tensor, target: Tensor, num_classes: int,
                           average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro", top_k: int = 1,
                           multidim_average: Literal["global", "samplewise"] = "global",
                           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target,
                           _multiclass_stat_scores_reports_values(top_k, average, multidim_average))
    if validate_args:
        _multiclass_stat_scores_arg_validation(num_classes, top_k, average, multidim_average, ignore_index)
        with fv:
            _multiclass_stat_scores_tensor_validation(preds, target, num_classes, multidim_average, ignore_index)
    tp, fp, tn, fn = _multiclass_stat_scores_update_fused(preds, target, num_classes, top_k, average,
                                                          multidim_average, ignore_index)
    fv.finish(preds.device, num_classes, errors=True)
    return _multiclass_stat_scores_compute(tp, fp, tn, fn, average, multidim_average)


# ======================================================================================= multilabel
def _multilabel_stat_scores_arg_validation(num_labels: int, threshold: float = 0.5, average: Optional[str] = "macro",
                                           multidim_average: str = "global", ignore_index: Optional[int] = None) -> None:
    C.check_num("num_labels", num_labels)
    C.check_threshold(threshold)
    C.check_choice("average", average, _AVG)
    C.check_choice("multidim_average", multidim_average, _MDA)
    C.check_ignore_index(ignore_index)


def _multilabel_stat_scores_tensor_validation(preds: Tensor, target: Tensor, num_labels: int,
                                              multidim_average: str, ignore_index: Optional[int] = None) -> None:
    C.check_multilabel_shapes(preds, target, num_labels)
    C.check_binary_values(target, "target", ignore_index)
    if not preds.is_floating_point():
        C.check_binary_values(preds, "preds")
    if multidim_average != "global" and preds.ndim < 3:
        raise ValueError("Expected input to be at least 3D when multidim_average is set to `samplewise`")


def _multilabel_stat_scores_format(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                                   ignore_index: Optional[int] = None) -> Tuple[Tensor, Tensor]:
    """Reference-contract formatting (stat_scores.py:647-669): (N, L, X) label tensors, ignored targets -1."""
    if preds.is_floating_point():
        if not C.in_unit_interval(preds):
            preds = preds.sigmoid()
        preds = preds > threshold
    preds = preds.reshape(*preds.shape[:2], -1)
    target = target.reshape(*target.shape[:2], -1)
    if ignore_index is not None:
        target = torch.where(target == ignore_index, torch.full_like(target, -1), target)
    return preds, target


def _multilabel_stat_scores_update(preds: Tensor, target: Tensor, multidim_average: str = "global"
                                   ) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """Per-label counts of formatted (N, L, X) tensors (stat_scores.py:672-681); global -> K2."""
    if multidim_average == "global":
        L = preds.shape[1]
        stats = torch.zeros(4, L, dtype=torch.int64, device=preds.device)
        _native.stat_scores_update_(stats, _label_major(preds, L), _label_major(target, L), L, 0.5, None,
                                    _native.PROB_AS_IS, report_values=False)
        return stats[0], stats[1], stats[2], stats[3]
    n, L = preds.shape[:2]
    rows = _native.stat_scores_rows(preds, target, preds.shape[2]).view(4, n, L)
    return rows[0].squeeze(), rows[1].squeeze(), rows[2].squeeze(), rows[3].squeeze()


def _multilabel_stat_scores_update_fused(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                                         multidim_average: str = "global", ignore_index: Optional[int] = None,
                                         out: Optional[Tensor] = None):
    """_format + _update in one K2 pass over the raw scores; `out` is an optional (4, L) int64 accumulator."""
    if multidim_average != "global":
        # rows = (sample, label) pairs of the (N, L, X) layout the inputs already have
        n = preds.shape[0]
        width = preds.numel() // (n * num_labels) if n else 0
        sigmoid = preds.is_floating_point() and not C.in_unit_interval(preds)
        rows = _native.stat_scores_rows(preds, target, width, threshold, ignore_index, sigmoid).view(4, n, num_labels)
        return rows[0].squeeze(), rows[1].squeeze(), rows[2].squeeze(), rows[3].squeeze()
    stats = out if out is not None else torch.zeros(4, num_labels, dtype=torch.int64, device=preds.device)
    _native.stat_scores_update_(stats, _label_major(preds, num_labels), _label_major(target, num_labels), num_labels,
                                threshold, ignore_index, _native.PROB_AUTO)
    return stats[0], stats[1], stats[2], stats[3]


def _multilabel_stat_scores_compute(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str] = "macro",
                                    multidim_average: str = "global") -> Tensor:
    res = torch.stack([tp, fp, tn, fn, tp + fn], dim=-1)
    sum_dim = 0 if multidim_average == "global" else 1
    if average == "micro":
        return res.sum(sum_dim)
    if average == "macro":
        return res.float().mean(sum_dim)
    if average == "weighted":
        w = tp + fn
        return (res * (w / w.sum()).reshape(*w.shape, 1)).sum(sum_dim)
    if average is None or average == "none":
        return res
    return None


def multilabel_stat_scores(preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                           average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                           multidim_average: Literal["global", "samplewise"] = "global",
                           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    fv = C.FusedValidation(validate_args, preds, target, multidim_average == "global")
    if validate_args:
        _multilabel_stat_scores_arg_validation(num_labels, threshold, average, multidim_average, ignore_index)
        with fv:
            _multilabel_stat_scores_tensor_validation(preds, target, num_labels, multidim_average, ignore_index)
    tp, fp, tn, fn = _multilabel_stat_scores_update_fused(preds, target, num_labels, threshold, multidim_average,
                                                          ignore_index)
    fv.finish(preds.device)
    return _multilabel_stat_scores_compute(tp, fp, tn, fn, average, multidim_average)


# ======================================================================================= legacy (Dice) path
def _del_column(data: Tensor, idx: int) -> Tensor:
    """stat_scores.py:820-822."""
    return torch.cat([data[:, :idx], data[:, (idx + 1):]], 1)


def _drop_negative_ignored_indices(preds: Tensor, target: Tensor, ignore_index: int, mode) -> Tuple[Tensor, Tensor]:
    """Remove samples whose target is a negative `ignore_index` before the case inference (stat_scores.py:825-858)."""
    from torchmetrics_b200.utilities.enums import DataType

    if mode == DataType.MULTIDIM_MULTICLASS and preds.dtype == torch.float:
        num_classes = preds.shape[1]
        preds = preds.transpose(1, preds.ndim - 1).reshape(-1, num_classes)
        target = target.reshape(-1)
    if mode in (DataType.MULTICLASS, DataType.MULTIDIM_MULTICLASS):
        keep = target != ignore_index
        preds, target = preds[keep], target[keep]
    return preds, target


def _stat_scores(preds: Tensor, target: Tensor, reduce: Optional[str] = "micro") -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """tp / fp / tn / fn of 0/1 tensors of shape (N, C) or (N, C, X) (stat_scores.py:861-906).

    The reference takes four masked sums over the dims `reduce` selects.  Every combination is either a column-wise
    count (K2: (N, C) with reduce="macro"), a whole-tensor count (K2 with one label) or a row-wise count (K2 row-wise
    form) of a contiguous view, so each runs as one pass of those kernels."""
    preds, target = preds.contiguous(), target.contiguous()
    dev = preds.device
    if preds.ndim == 2:
        n, c = preds.shape
        if reduce == "micro":
            stats = torch.zeros(4, 1, dtype=torch.int64, device=dev)
            _native.stat_scores_update_(stats, preds, target, 1, 0.5, None, _native.PROB_AS_IS, report_values=False)
            return stats[0, 0], stats[1, 0], stats[2, 0], stats[3, 0]
        if reduce == "macro":
            stats = torch.zeros(4, c, dtype=torch.int64, device=dev)
            if n:
                _native.stat_scores_update_(stats, preds, target, c, 0.5, None, _native.PROB_AS_IS,
                                            report_values=False)
            return stats[0], stats[1], stats[2], stats[3]
        rows = _native.stat_scores_rows(preds, target, c)  # "samples": one row per sample
        return rows[0], rows[1], rows[2], rows[3]
    n, c, x = preds.shape
    if reduce == "micro":
        rows = _native.stat_scores_rows(preds, target, c * x)
        return rows[0], rows[1], rows[2], rows[3]
    if reduce == "macro":
        rows = _native.stat_scores_rows(preds, target, x).view(4, n, c)
        return rows[0], rows[1], rows[2], rows[3]
    # "samples": the class dimension is summed out -> make it the fastest one
    rows = _native.stat_scores_rows(preds.transpose(1, 2).contiguous(), target.transpose(1, 2).contiguous(), c)
    rows = rows.view(4, n, x)
    return rows[0], rows[1], rows[2], rows[3]


def _stat_scores_update(preds: Tensor, target: Tensor, reduce: Optional[str] = "micro",
                        mdmc_reduce: Optional[str] = None, num_classes: Optional[int] = None,
                        top_k: Optional[int] = 1, threshold: float = 0.5, multiclass: Optional[bool] = None,
                        ignore_index: Optional[int] = None, mode=None) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    """The legacy update behind Dice (stat_scores.py:909-995): case inference, then `_stat_scores`."""
    from torchmetrics_b200.utilities.checks import _input_format_classification

    negative_dropped = False
    if ignore_index is not None and ignore_index < 0 and mode is not None:
        preds, target = _drop_negative_ignored_indices(preds, target, ignore_index, mode)
        negative_dropped = True
    preds, target, _ = _input_format_classification(preds, target, threshold=threshold, num_classes=num_classes,
                                                    multiclass=multiclass, top_k=top_k, ignore_index=ignore_index)
    if ignore_index is not None and ignore_index >= preds.shape[1]:
        raise ValueError(f"The `ignore_index` {ignore_index} is not valid for inputs with {preds.shape[1]} classes")
    if ignore_index is not None and preds.shape[1] == 1:
        raise ValueError("You can not use `ignore_index` with binary data.")
    if preds.ndim == 3:
        if not mdmc_reduce:
            raise ValueError(
                "When your inputs are multi-dimensional multi-class, you have to set the `mdmc_reduce` parameter"
            )
        if mdmc_reduce == "global":
            preds = torch.transpose(preds, 1, 2).reshape(-1, preds.shape[1])
            target = torch.transpose(target, 1, 2).reshape(-1, target.shape[1])
    drop_column = ignore_index is not None and not negative_dropped
    if drop_column and reduce != "macro":
        preds, target = _del_column(preds, ignore_index), _del_column(target, ignore_index)
    tp, fp, tn, fn = _stat_scores(preds, target, reduce=reduce)
    if drop_column and reduce == "macro":
        for stat in (tp, fp, tn, fn):
            stat[..., ignore_index] = -1
    return tp, fp, tn, fn


def _stat_scores_compute(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor) -> Tensor:
    """[tp, fp, tn, fn, support] along a new last dim, negatives (ignored classes) kept at -1 (:998-1018)."""
    out = torch.stack([tp, fp, tn, fn, tp + fn], dim=-1)
    return torch.where(out < 0, torch.full_like(out, -1), out)


def _reduce_stat_scores(numerator: Tensor, denominator: Tensor, weights: Optional[Tensor], average: Optional[str],
                        mdmc_average: Optional[str], zero_division: int = 0) -> Tensor:
    """numerator / denominator (optionally weighted) with ignored (< 0) and empty (== 0) denominators handled
    (stat_scores.py:1021-1075)."""
    numerator, denominator = numerator.float(), denominator.float()
    empty = denominator == 0
    ignored = denominator < 0
    weights = torch.ones_like(denominator) if weights is None else weights.float()
    numerator = torch.where(empty, torch.full_like(numerator, float(zero_division)), numerator)
    denominator = torch.where(empty | ignored, torch.ones_like(denominator), denominator)
    weights = torch.where(ignored, torch.zeros_like(weights), weights)
    if average not in ("micro", "none", None):
        weights = weights / weights.sum(dim=-1, keepdim=True)
    scores = weights * (numerator / denominator)
    # sum(weights) == 0: the only class present was ignored under average="weighted"
    scores = torch.where(torch.isnan(scores), torch.full_like(scores, float(zero_division)), scores)
    if mdmc_average == "samplewise":
        scores = scores.mean(dim=0)
        ignored = ignored.sum(dim=0).bool()
    if average in ("none", None):
        return torch.where(ignored, torch.full_like(scores, float("nan")), scores)
    return scores.sum()


# ======================================================================================= task wrapper
def stat_scores(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"],
                threshold: float = 0.5, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
                multidim_average: Optional[Literal["global", "samplewise"]] = "global", top_k: Optional[int] = 1,
                ignore_index: Optional[int] = None, validate_args: bool = True) -> Tensor:
    task = ClassificationTask.from_str(task)
    assert multidim_average is not None
    if task == ClassificationTask.BINARY:
        return binary_stat_scores(preds, target, threshold, multidim_average, ignore_index, validate_args)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        if not isinstance(top_k, int):
            raise ValueError(f"`top_k` is expected to be `int` but `{type(top_k)} was passed.`")
        return multiclass_stat_scores(preds, target, num_classes, average, top_k, multidim_average, ignore_index,
                                      validate_args)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_stat_scores(preds, target, num_labels, threshold, average, multidim_average, ignore_index,
                                      validate_args)
    raise ValueError(f"Task {task} not supported.")
