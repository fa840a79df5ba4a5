"""Group fairness (reference: functional/classification/group_fairness.py).

The reference sorts the samples by group, splits them and calls `_binary_stat_scores_update` once per group
(:52-84: a sort, a host sync for the split sizes and 4 reductions per group).  The four counts of every group are the
cells of a (group, 2*target + pred) contingency table, so here ONE pass of the confusion-matrix kernel (K1) over the
formatted labels produces all groups at once: rows = group id, columns 0..3 = tn / fp / fn / tp, column 4 = ignored.
"""
from typing import Dict, List, Optional, Tuple

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200 import _native
from torchmetrics_b200.functional.classification.stat_scores import (
    _binary_stat_scores_arg_validation,
    _binary_stat_scores_format,
    _binary_stat_scores_tensor_validation,
)
from torchmetrics_b200.utilities.compute import _safe_divide
from torchmetrics_b200.utilities.prints import rank_zero_warn

_Stats = Tuple[Tensor, Tensor, Tensor, Tensor]


def _groups_validation(groups: Tensor, num_groups: int) -> None:
    if torch.max(groups) > num_groups:
        raise ValueError(
            f"The largest number in the groups tensor is {torch.max(groups)}, which is larger than the specified",
            f"number of groups {num_groups}. The group identifiers should be ``0, 1, ..., (num_groups - 1)``.",
        )
    if groups.dtype != torch.long:
        raise ValueError(f"Expected dtype of argument groups to be long, not {groups.dtype}.")


def _groups_format(groups: Tensor) -> Tensor:
    return groups.reshape(groups.shape[0], -1)


def _group_table(preds: Tensor, target: Tensor, groups: Tensor, num_ids: int) -> Tensor:
    """(num_ids, 5) int64 [tn, fp, fn, tp, ignored] per group id from formatted labels (target -1 = ignored)."""
    side = max(int(num_ids), 5)
    t = target.reshape(-1).long()
    col = torch.where(t < 0, torch.full_like(t, 4), 2 * t + preds.reshape(-1).long())
    table = torch.zeros(side, side, dtype=torch.int64, device=preds.device)
    _native.confmat_update_(table, col, groups.reshape(-1), side)
    _native.raise_on_status(preds.device, side)
    return table[:num_ids, :5]


def _binary_groups_stat_scores(preds: Tensor, target: Tensor, groups: Tensor, num_groups: int, threshold: float = 0.5,
                               ignore_index: Optional[int] = None, validate_args: bool = True) -> List[_Stats]:
    """(tp, fp, tn, fn) of every group that occurs, in ascending group id (group_fairness.py:52-84)."""
    if validate_args:
        _binary_stat_scores_arg_validation(threshold, "global", ignore_index)
        _binary_stat_scores_tensor_validation(preds, target, "global", ignore_index)
        _groups_validation(groups, num_groups)
    preds, target = _binary_stat_scores_format(preds, target, threshold, ignore_index)
    groups = _groups_format(groups).to(preds.device)
    num_ids = max(int(groups.max()) + 1, 1) if groups.numel() else 0
    table = _group_table(preds, target.to(preds.device), groups, num_ids)
    present = torch.nonzero(table.sum(1) > 0).flatten().tolist()  # the reference's split sizes need the same sync
    return [(table[g, 3], table[g, 1], table[g, 0], table[g, 2]) for g in present]


def _groups_reduce(group_stats: List[_Stats]) -> Dict[str, Tensor]:
    return {f"group_{g}": torch.stack(s) / torch.stack(s).sum() for g, s in enumerate(group_stats)}


def _groups_stat_transform(group_stats: List[_Stats]) -> Dict[str, Tensor]:
    return {name: torch.stack([s[i] for s in group_stats]) for i, name in enumerate(("tp", "fp", "tn", "fn"))}


def binary_groups_stat_rates(preds: Tensor, target: Tensor, groups: Tensor, num_groups: int, threshold: float = 0.5,
                             ignore_index: Optional[int] = None, validate_args: bool = True) -> Dict[str, Tensor]:
    return _groups_reduce(_binary_groups_stat_scores(preds, target, groups, num_groups, threshold, ignore_index,
                                                     validate_args))


def _rate_ratio(prefix: str, rates: Tensor) -> Dict[str, Tensor]:
    lo, hi = torch.argmin(rates), torch.argmax(rates)
    return {f"{prefix}_{lo}_{hi}": _safe_divide(rates[lo], rates[hi])}


def _compute_binary_demographic_parity(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor) -> Dict[str, Tensor]:
    return _rate_ratio("DP", _safe_divide(tp + fp, tp + fp + tn + fn))


def _compute_binary_equal_opportunity(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor) -> Dict[str, Tensor]:
    return _rate_ratio("EO", _safe_divide(tp, tp + fn))


def demographic_parity(preds: Tensor, groups: Tensor, threshold: float = 0.5, ignore_index: Optional[int] = None,
                       validate_args: bool = True) -> Dict[str, Tensor]:
    num_groups = torch.unique(groups).shape[0]
    target = torch.zeros(preds.shape, device=preds.device)
    stats = _binary_groups_stat_scores(preds, target, groups, num_groups, threshold, ignore_index, validate_args)
    return _compute_binary_demographic_parity(**_groups_stat_transform(stats))


def equal_opportunity(preds: Tensor, target: Tensor, groups: Tensor, threshold: float = 0.5,
                      ignore_index: Optional[int] = None, validate_args: bool = True) -> Dict[str, Tensor]:
    num_groups = torch.unique(groups).shape[0]
    stats = _binary_groups_stat_scores(preds, target, groups, num_groups, threshold, ignore_index, validate_args)
    return _compute_binary_equal_opportunity(**_groups_stat_transform(stats))


def binary_fairness(preds: Tensor, target: Tensor, groups: Tensor,
                    task: Literal["demographic_parity", "equal_opportunity", "all"] = "all", threshold: float = 0.5,
                    ignore_index: Optional[int] = None, validate_args: bool = True) -> Dict[str, Tensor]:
    if task not in ["demographic_parity", "equal_opportunity", "all"]:
        raise ValueError(
            f"Expected argument `task` to either be ``demographic_parity``,"
            f"``equal_opportunity`` or ``all`` but got {task}."
        )
    if task == "demographic_parity":
        if target is not None:
            rank_zero_warn("The task demographic_parity does not require a target.", UserWarning)
        target = torch.zeros(preds.shape, device=preds.device)
    num_groups = torch.unique(groups).shape[0]
    stats = _groups_stat_transform(
        _binary_groups_stat_scores(preds, target, groups, num_groups, threshold, ignore_index, validate_args))
    out: Dict[str, Tensor] = {}
    if task in ("demographic_parity", "all"):
        out.update(_compute_binary_demographic_parity(**stats))
    if task in ("equal_opportunity", "all"):
        out.update(_compute_binary_equal_opportunity(**stats))
    return out
