"""Stateful Dice (reference: classification/dice.py): the one consumer of the legacy `_stat_scores_update`."""
from typing import Any, Optional, Tuple

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.functional.classification.dice import _dice_compute
from torchmetrics_b200.functional.classification.stat_scores import _stat_scores_update
from torchmetrics_b200.metric import Metric


class Dice(Metric):
    is_differentiable = False
    higher_is_better = True
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0
    plot_legend_name = "Class"

    def __init__(self, zero_division: int = 0, num_classes: Optional[int] = None, threshold: float = 0.5,
                 average: Optional[Literal["micro", "macro", "none"]] = "micro",
                 mdmc_average: Optional[str] = "global", ignore_index: Optional[int] = None,
                 top_k: Optional[int] = None, multiclass: Optional[bool] = None, **kwargs: Any) -> None:
        kwargs.pop("reduce", None)
        kwargs.pop("mdmc_reduce", None)
        super().__init__(**kwargs)
        allowed_average = ("micro", "macro", "samples", "none", None)
        if average not in allowed_average:
            raise ValueError(f"The `average` has to be one of {allowed_average}, got {average}.")
        if average not in ["micro", "macro", "samples"]:
            raise ValueError(f"The `reduce` {average} is not valid.")
        if mdmc_average not in [None, "samplewise", "global"]:
            raise ValueError(f"The `mdmc_reduce` {mdmc_average} is not valid.")
        if average == "macro" and (not num_classes or num_classes < 1):
            raise ValueError("When you set `average` as 'macro', you have to provide the number of classes.")
        if num_classes and ignore_index is not None and (not ignore_index < num_classes or num_classes == 1):
            raise ValueError(f"The `ignore_index` {ignore_index} is not valid for inputs with {num_classes} classes")
        self.reduce = average
        self.mdmc_reduce = mdmc_average
        self.num_classes = num_classes
        self.threshold = threshold
        self.multiclass = multiclass
        self.ignore_index = ignore_index
        self.top_k = top_k
        self.average = average
        self.zero_division = zero_division
        # per-sample statistics are cat states, everything else constant-size sums
        self._per_sample = mdmc_average == "samplewise" or average == "samples"
        for name in ("tp", "fp", "tn", "fn"):
            if self._per_sample:
                self.add_state(name, default=[], dist_reduce_fx="cat")
            else:
                shape = [] if average == "micro" else [num_classes]
                self.add_state(name, default=torch.zeros(shape, dtype=torch.long), dist_reduce_fx="sum")

    def update(self, preds: Tensor, target: Tensor) -> None:
        stats = _stat_scores_update(preds, target, reduce=self.reduce, mdmc_reduce=self.mdmc_reduce,
                                    threshold=self.threshold, num_classes=self.num_classes, top_k=self.top_k,
                                    multiclass=self.multiclass, ignore_index=self.ignore_index)
        for name, value in zip(("tp", "fp", "tn", "fn"), stats):
            if self._per_sample:
                getattr(self, name).append(value)
            else:
                setattr(self, name, getattr(self, name) + value)

    def _get_final_stats(self) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
        return tuple(torch.cat(s) if isinstance(s, list) else s for s in (self.tp, self.fp, self.tn, self.fn))

    def compute(self) -> Tensor:
        tp, fp, _, fn = self._get_final_stats()
        return _dice_compute(tp, fp, fn, self.average, self.mdmc_reduce, self.zero_division)
