"""Shared pieces of the nominal association measures (reference: functional/nominal/utils.py)."""
from typing import Optional, Tuple

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.utilities.prints import rank_zero_warn


def _nominal_input_validation(nan_strategy: str, nan_replace_value: Optional[float]) -> None:
    if nan_strategy not in ["replace", "drop"]:
        raise ValueError(
            f"Argument `nan_strategy` is expected to be one of `['replace', 'drop']`, but got {nan_strategy}"
        )
    if nan_strategy == "replace" and not isinstance(nan_replace_value, (float, int)):
        raise ValueError(
            "Argument `nan_replace` is expected to be of a type `int` or `float` when `nan_strategy = 'replace`, "
            f"but got {nan_replace_value}"
        )


def _handle_nan_in_data(preds: Tensor, target: Tensor, nan_strategy: Literal["replace", "drop"] = "replace",
                        nan_replace_value: Optional[float] = 0.0) -> Tuple[Tensor, Tensor]:
    if nan_strategy == "replace":
        return preds.nan_to_num(nan_replace_value), target.nan_to_num(nan_replace_value)
    bad = torch.logical_or(preds.isnan(), target.isnan())
    return preds[~bad], target[~bad]


def _contingency_update(preds: Tensor, target: Tensor, num_classes: int, nan_strategy: str,
                        nan_replace_value: Optional[float]) -> Tensor:
    """The `_*_update` all four measures share (e.g. cramers.py:31-57): argmax of 2-D inputs, NaN policy, then the
    multiclass confusion matrix - one K1 pass, exact int64 counts."""
    from torchmetrics_b200.functional.classification.confusion_matrix import _multiclass_confusion_matrix_update

    preds = preds.argmax(1) if preds.ndim == 2 else preds
    target = target.argmax(1) if target.ndim == 2 else target
    preds, target = _handle_nan_in_data(preds, target, nan_strategy, nan_replace_value)
    return _multiclass_confusion_matrix_update(preds, target, num_classes)


def _compute_expected_freqs(confmat: Tensor) -> Tensor:
    rows, cols = confmat.sum(1), confmat.sum(0)
    return torch.einsum("r, c -> rc", rows, cols) / confmat.sum()


def _compute_chi_squared(confmat: Tensor, bias_correction: bool) -> Tensor:
    """Pearson chi-squared with Yates' correction for one degree of freedom (utils.py:46-67)."""
    expected = _compute_expected_freqs(confmat)
    df = expected.numel() - sum(expected.shape) + expected.ndim - 1
    if df == 0:
        return torch.tensor(0.0, device=confmat.device)
    if df == 1 and bias_correction:
        confmat = confmat.to(expected.dtype)
        direction = (expected - confmat).sign()
        confmat = confmat + direction * torch.minimum(0.5 * torch.ones_like(direction), direction.abs())
    return torch.sum((confmat - expected) ** 2 / expected)


def _drop_empty_rows_and_cols(confmat: Tensor) -> Tensor:
    confmat = confmat[confmat.sum(1) != 0]
    return confmat[:, confmat.sum(0) != 0]


def _compute_phi_squared_corrected(phi_squared: Tensor, num_rows: int, num_cols: int, confmat_sum: Tensor) -> Tensor:
    return torch.max(torch.tensor(0.0, device=phi_squared.device),
                     phi_squared - ((num_rows - 1) * (num_cols - 1)) / (confmat_sum - 1))


def _compute_rows_and_cols_corrected(num_rows: int, num_cols: int, confmat_sum: Tensor) -> Tuple[Tensor, Tensor]:
    return (num_rows - (num_rows - 1) ** 2 / (confmat_sum - 1), num_cols - (num_cols - 1) ** 2 / (confmat_sum - 1))


def _compute_bias_corrected_values(phi_squared: Tensor, num_rows: int, num_cols: int, confmat_sum: Tensor):
    rows, cols = _compute_rows_and_cols_corrected(num_rows, num_cols, confmat_sum)
    return _compute_phi_squared_corrected(phi_squared, num_rows, num_cols, confmat_sum), rows, cols


def _unable_to_use_bias_correction_warning(metric_name: str) -> None:
    rank_zero_warn(
        f"Unable to compute {metric_name} using bias correction. Please consider to set `bias_correction=False`."
    )


def _pairwise_matrix(matrix: Tensor, update, compute, symmetric: bool = True) -> Tensor:
    """The `*_matrix` functions: the measure for every pair of columns of an (N, V) table of categories."""
    import itertools

    n_var = matrix.shape[1]
    out = torch.ones(n_var, n_var, device=matrix.device)
    for i, j in itertools.combinations(range(n_var), 2):
        x, y = matrix[:, i], matrix[:, j]
        confmat = update(x, y, len(torch.cat([x, y]).unique()))
        out[i, j] = compute(confmat)
        out[j, i] = out[i, j] if symmetric else compute(confmat.T)
    return out
