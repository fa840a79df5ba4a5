"""Stateful group fairness metrics (reference: classification/group_fairness.py); one K1 pass per update."""
from typing import Any, Dict, List, Optional, Tuple

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.functional.classification.group_fairness import (
    _binary_groups_stat_scores,
    _compute_binary_demographic_parity,
    _compute_binary_equal_opportunity,
)
from torchmetrics_b200.functional.classification.stat_scores import _binary_stat_scores_arg_validation
from torchmetrics_b200.metric import Metric
from torchmetrics_b200.utilities.prints import rank_zero_warn


class _AbstractGroupStatScores(Metric):
    tp: Tensor
    fp: Tensor
    tn: Tensor
    fn: Tensor

    def _create_states(self, num_groups: int) -> None:
        for name in ("tp", "fp", "tn", "fn"):
            self.add_state(name, torch.zeros(num_groups, dtype=torch.long), dist_reduce_fx="sum")

    def _update_states(self, group_stats: List[Tuple[Tensor, Tensor, Tensor, Tensor]]) -> None:
        # the reference indexes the states by position in the list of groups that occurred (:55-61)
        if not group_stats:
            return
        k = len(group_stats)
        for i, name in enumerate(("tp", "fp", "tn", "fn")):
            getattr(self, name)[:k] += torch.stack([s[i] for s in group_stats])

    def _init_common(self, num_groups: int, threshold: float, ignore_index: Optional[int], validate_args: bool) -> None:
        if validate_args:
            _binary_stat_scores_arg_validation(threshold, "global", ignore_index)
        if not isinstance(num_groups, int) and num_groups < 2:
            raise ValueError(f"Expected argument `num_groups` to be an int larger than 1, but got {num_groups}")
        self.num_groups = num_groups
        self.threshold = threshold
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._create_states(num_groups)


class BinaryGroupStatRates(_AbstractGroupStatScores):
    is_differentiable = False
    higher_is_better = False
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0

    def __init__(self, num_groups: int, threshold: float = 0.5, ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__()
        self._init_common(num_groups, threshold, ignore_index, validate_args)

    def update(self, preds: Tensor, target: Tensor, groups: Tensor) -> None:
        self._update_states(_binary_groups_stat_scores(preds, target, groups, self.num_groups, self.threshold,
                                                       self.ignore_index, self.validate_args))

    def compute(self) -> Dict[str, Tensor]:
        results = torch.stack((self.tp, self.fp, self.tn, self.fn), dim=1)
        return {f"group_{i}": group / group.sum() for i, group in enumerate(results)}


class BinaryFairness(_AbstractGroupStatScores):
    is_differentiable = False
    higher_is_better = False
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0

    def __init__(self, num_groups: int, task: Literal["demographic_parity", "equal_opportunity", "all"] = "all",
                 threshold: float = 0.5, ignore_index: Optional[int] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__()
        if task not in ["demographic_parity", "equal_opportunity", "all"]:
            raise ValueError(
                f"Expected argument `task` to either be ``demographic_parity``,"
                f"``equal_opportunity`` or ``all`` but got {task}."
            )
        self.task = task
        self._init_common(num_groups, threshold, ignore_index, validate_args)

    def update(self, preds: Tensor, target: Tensor, groups: Tensor) -> None:
        if self.task == "demographic_parity":
            if target is not None:
                rank_zero_warn("The task demographic_parity does not require a target.", UserWarning)
            target = torch.zeros(preds.shape, device=preds.device)
        self._update_states(_binary_groups_stat_scores(preds, target, groups, self.num_groups, self.threshold,
                                                       self.ignore_index, self.validate_args))

    def compute(self) -> Dict[str, Tensor]:
        out: Dict[str, Tensor] = {}
        if self.task in ("demographic_parity", "all"):
            out.update(_compute_binary_demographic_parity(self.tp, self.fp, self.tn, self.fn))
        if self.task in ("equal_opportunity", "all"):
            out.update(_compute_binary_equal_opportunity(self.tp, self.fp, self.tn, self.fn))
        return out
