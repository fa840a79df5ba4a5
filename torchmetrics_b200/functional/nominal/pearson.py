"""Pearson's contingency coefficient (reference: functional/nominal/pearson.py)."""
from typing import Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.functional.nominal import utils as U


def _pearsons_contingency_coefficient_update(preds: Tensor, target: Tensor, num_classes: int,
                                             nan_strategy: Literal["replace", "drop"] = "replace",
                                             nan_replace_value: Optional[float] = 0.0) -> Tensor:
    return U._contingency_update(preds, target, num_classes, nan_strategy, nan_replace_value)


def _pearsons_contingency_coefficient_compute(confmat: Tensor) -> Tensor:
    """pearson.py:57-72: sqrt(phi^2 / (1 + phi^2))."""
    confmat = U._drop_empty_rows_and_cols(confmat)
    phi2 = U._compute_chi_squared(confmat, bias_correction=False) / confmat.sum()
    return torch.sqrt(phi2 / (1 + phi2)).clamp(0.0, 1.0)


def pearsons_contingency_coefficient(preds: Tensor, target: Tensor,
                                     nan_strategy: Literal["replace", "drop"] = "replace",
                                     nan_replace_value: Optional[float] = 0.0) -> Tensor:
    U._nominal_input_validation(nan_strategy, nan_replace_value)
    num_classes = len(torch.cat([preds, target]).unique())
    return _pearsons_contingency_coefficient_compute(
        _pearsons_contingency_coefficient_update(preds, target, num_classes, nan_strategy, nan_replace_value))


def pearsons_contingency_coefficient_matrix(matrix: Tensor, nan_strategy: Literal["replace", "drop"] = "replace",
                                            nan_replace_value: Optional[float] = 0.0) -> Tensor:
    U._nominal_input_validation(nan_strategy, nan_replace_value)
    return U._pairwise_matrix(
        matrix, lambda x, y, c: _pearsons_contingency_coefficient_update(x, y, c, nan_strategy, nan_replace_value),
        _pearsons_contingency_coefficient_compute)
