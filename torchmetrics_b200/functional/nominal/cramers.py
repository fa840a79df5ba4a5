"""Cramer's V (reference: functional/nominal/cramers.py)."""
from typing import Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.functional.nominal import utils as U


def _cramers_v_update(preds: Tensor, target: Tensor, num_classes: int,
                      nan_strategy: Literal["replace", "drop"] = "replace",
                      nan_replace_value: Optional[float] = 0.0) -> Tensor:
    return U._contingency_update(preds, target, num_classes, nan_strategy, nan_replace_value)


def _cramers_v_compute(confmat: Tensor, bias_correction: bool) -> Tensor:
    """cramers.py:60-89."""
    confmat = U._drop_empty_rows_and_cols(confmat)
    total = confmat.sum()
    phi2 = U._compute_chi_squared(confmat, bias_correction) / total
    n_rows, n_cols = confmat.shape
    if bias_correction:
        phi2c, rows_c, cols_c = U._compute_bias_corrected_values(phi2, n_rows, n_cols, total)
        if torch.min(rows_c, cols_c) == 1:
            U._unable_to_use_bias_correction_warning(metric_name="Cramer's V")
            return torch.tensor(float("nan"), device=confmat.device)
        value = torch.sqrt(phi2c / torch.min(rows_c - 1, cols_c - 1))
    else:
        value = torch.sqrt(phi2 / min(n_rows - 1, n_cols - 1))
    return value.clamp(0.0, 1.0)


def cramers_v(preds: Tensor, target: Tensor, bias_correction: bool = True,
              nan_strategy: Literal["replace", "drop"] = "replace", nan_replace_value: Optional[float] = 0.0) -> Tensor:
    U._nominal_input_validation(nan_strategy, nan_replace_value)
    num_classes = len(torch.cat([preds, target]).unique())
    return _cramers_v_compute(_cramers_v_update(preds, target, num_classes, nan_strategy, nan_replace_value),
                              bias_correction)


def cramers_v_matrix(matrix: Tensor, bias_correction: bool = True,
                     nan_strategy: Literal["replace", "drop"] = "replace",
                     nan_replace_value: Optional[float] = 0.0) -> Tensor:
    U._nominal_input_validation(nan_strategy, nan_replace_value)
    return U._pairwise_matrix(matrix, lambda x, y, c: _cramers_v_update(x, y, c, nan_strategy, nan_replace_value),
                              lambda cm: _cramers_v_compute(cm, bias_correction))
