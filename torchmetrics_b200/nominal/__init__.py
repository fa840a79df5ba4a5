"""Stateful nominal association measures (reference: nominal/{cramers,tschuprows,pearson,theils_u}.py).

State = the (C, C) contingency table, accumulated in place by the K1 kernel as exact int64 counts (the reference keeps
a float32 table and adds a fresh matrix to it per update) and sum-reduced across ranks; `compute` is the O(C^2)
reducer of the functional module."""
from typing import Any, Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200 import _native
from importlib import import_module

# the functional package re-exports functions named like its sub-modules, so fetch the modules explicitly
_cr = import_module("torchmetrics_b200.functional.nominal.cramers")
_pe = import_module("torchmetrics_b200.functional.nominal.pearson")
_th = import_module("torchmetrics_b200.functional.nominal.theils_u")
_ts = import_module("torchmetrics_b200.functional.nominal.tschuprows")
from torchmetrics_b200.functional.nominal.utils import _handle_nan_in_data, _nominal_input_validation
from torchmetrics_b200.metric import Metric


class _NominalBase(Metric):
    full_state_update = False
    is_differentiable = False
    higher_is_better = True
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0
    confmat: Tensor

    def _init_nominal(self, num_classes: int, nan_strategy: str, nan_replace_value: Optional[float]) -> None:
        self.num_classes = num_classes
        _nominal_input_validation(nan_strategy, nan_replace_value)
        self.nan_strategy = nan_strategy
        self.nan_replace_value = nan_replace_value
        self.add_state("confmat", torch.zeros(num_classes, num_classes, dtype=torch.long), dist_reduce_fx="sum")

    def update(self, preds: Tensor, target: Tensor) -> None:
        preds = preds.argmax(1) if preds.ndim == 2 else preds
        target = target.argmax(1) if target.ndim == 2 else target
        preds, target = _handle_nan_in_data(preds, target, self.nan_strategy, self.nan_replace_value)
        _native.confmat_update_(self.confmat, preds, target, self.num_classes)


class CramersV(_NominalBase):
    def __init__(self, num_classes: int, bias_correction: bool = True,
                 nan_strategy: Literal["replace", "drop"] = "replace", nan_replace_value: Optional[float] = 0.0,
                 **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self.bias_correction = bias_correction
        self._init_nominal(num_classes, nan_strategy, nan_replace_value)

    def compute(self) -> Tensor:
        _native.raise_on_status(self.confmat.device, self.num_classes) if self.confmat.is_cuda else None
        return _cr._cramers_v_compute(self.confmat, self.bias_correction)


class TschuprowsT(_NominalBase):
    def __init__(self, num_classes: int, bias_correction: bool = True,
                 nan_strategy: Literal["replace", "drop"] = "replace", nan_replace_value: Optional[float] = 0.0,
                 **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self.bias_correction = bias_correction
        self._init_nominal(num_classes, nan_strategy, nan_replace_value)

    def compute(self) -> Tensor:
        _native.raise_on_status(self.confmat.device, self.num_classes) if self.confmat.is_cuda else None
        return _ts._tschuprows_t_compute(self.confmat, self.bias_correction)


class PearsonsContingencyCoefficient(_NominalBase):
    def __init__(self, num_classes: int, nan_strategy: Literal["replace", "drop"] = "replace",
                 nan_replace_value: Optional[float] = 0.0, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._init_nominal(num_classes, nan_strategy, nan_replace_value)

    def compute(self) -> Tensor:
        _native.raise_on_status(self.confmat.device, self.num_classes) if self.confmat.is_cuda else None
        return _pe._pearsons_contingency_coefficient_compute(self.confmat)


class TheilsU(_NominalBase):
    def __init__(self, num_classes: int, nan_strategy: Literal["replace", "drop"] = "replace",
                 nan_replace_value: Optional[float] = 0.0, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        self._init_nominal(num_classes, nan_strategy, nan_replace_value)

    def compute(self) -> Tensor:
        _native.raise_on_status(self.confmat.device, self.num_classes) if self.confmat.is_cuda else None
        return _th._theils_u_compute(self.confmat)


__all__ = ["CramersV", "PearsonsContingencyCoefficient", "TheilsU", "TschuprowsT"]
