"""Theil's U, the uncertainty coefficient (reference: functional/nominal/theils_u.py)."""
from typing import Optional

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.functional.nominal import utils as U


def _conditional_entropy_compute(confmat: Tensor) -> Tensor:
    """H(X | Y) of the table's joint distribution (theils_u.py:29-49)."""
    confmat = U._drop_empty_rows_and_cols(confmat)
    total = confmat.sum()
    p_xy = confmat / total
    p_y = (confmat.sum(1) / total).unsqueeze(1).repeat(1, p_xy.shape[1])
    return torch.nansum(p_xy * torch.log(p_y / p_xy))


def _theils_u_update(preds: Tensor, target: Tensor, num_classes: int,
                     nan_strategy: Literal["replace", "drop"] = "replace",
                     nan_replace_value: Optional[float] = 0.0) -> Tensor:
    return U._contingency_update(preds, target, num_classes, nan_strategy, nan_replace_value)


def _theils_u_compute(confmat: Tensor) -> Tensor:
    """theils_u.py:80-103."""
    confmat = U._drop_empty_rows_and_cols(confmat)
    s_xy = _conditional_entropy_compute(confmat)
    p_x = confmat.sum(0) / confmat.sum()
    s_x = -torch.sum(p_x * torch.log(p_x))
    if s_x == 0:
        return torch.tensor(0, device=confmat.device)
    return (s_x - s_xy) / s_x


def theils_u(preds: Tensor, target: Tensor, nan_strategy: Literal["replace", "drop"] = "replace",
             nan_replace_value: Optional[float] = 0.0) -> Tensor:
    num_classes = len(torch.cat([preds, target]).unique())
    return _theils_u_compute(_theils_u_update(preds, target, num_classes, nan_strategy, nan_replace_value))


def theils_u_matrix(matrix: Tensor, nan_strategy: Literal["replace", "drop"] = "replace",
                    nan_replace_value: Optional[float] = 0.0) -> Tensor:
    U._nominal_input_validation(nan_strategy, nan_replace_value)
    return U._pairwise_matrix(matrix, lambda x, y, c: _theils_u_update(x, y, c, nan_strategy, nan_replace_value),
                              _theils_u_compute, symmetric=False)
