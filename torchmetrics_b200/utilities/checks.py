"""Input checks on the hot path (reference: utilities/checks.py:39-44)."""
from torch import Tensor


def _check_same_shape(preds: Tensor, target: Tensor) -> None:
    if preds.shape != target.shape:
        raise RuntimeError(
            f"Predictions and targets are expected to have the same shape, but got {preds.shape} and {target.shape}."
        )


# ------------------------------------------------------------------------------------------------------------------
# Legacy input inference of `_stat_scores_update` / Dice (reference: utilities/checks.py:39-456).  Host logic only:
# it decides which of four input cases a (preds, target) pair is, validates it and brings both tensors to 0/1 int
# tensors of shape (N, C) or (N, C, X).  The value probes (min / max of both tensors) are read once, in one
# device-to-host copy, instead of once per rule.
# ------------------------------------------------------------------------------------------------------------------
from typing import Optional, Tuple  # noqa: E402

import torch  # noqa: E402

from torchmetrics_b200.utilities.data import select_topk, to_onehot  # noqa: E402
from torchmetrics_b200.utilities.enums import DataType  # noqa: E402


def _check_for_empty_tensors(preds: Tensor, target: Tensor) -> bool:
    return preds.numel() == target.numel() == 0


class _Extremes:
    """min / max of preds and target, fetched together on first use (one sync for all the rules below)."""

    def __init__(self, preds: Tensor, target: Tensor) -> None:
        self._p, self._t, self._v = preds, target, None

    def _get(self):
        if self._v is None:
            parts = []
            for x in (self._p, self._t):
                if x.numel():
                    lo, hi = torch.aminmax(x)
                    parts += [lo.double(), hi.double()]
                else:
                    z = torch.zeros((), dtype=torch.double, device=x.device)
                    parts += [z, z]
            self._v = torch.stack([p.to(parts[0].device) for p in parts]).tolist()
        return self._v

    pmin = property(lambda self: self._get()[0])
    pmax = property(lambda self: self._get()[1])
    tmin = property(lambda self: self._get()[2])
    tmax = property(lambda self: self._get()[3])


def _basic_input_validation(preds: Tensor, target: Tensor, threshold: float, multiclass: Optional[bool],
                            ignore_index: Optional[int], ext: Optional[_Extremes] = None) -> None:
    """checks.py:47-76."""
    if _check_for_empty_tensors(preds, target):
        return
    ext = ext or _Extremes(preds, target)
    if target.is_floating_point():
        raise ValueError("The `target` has to be an integer tensor.")
    if (ignore_index is None and ext.tmin < 0) or (ignore_index and ignore_index >= 0 and ext.tmin < 0):
        raise ValueError("The `target` has to be a non-negative tensor.")
    preds_float = preds.is_floating_point()
    if not preds_float and ext.pmin < 0:
        raise ValueError("If `preds` are integers, they have to be non-negative.")
    if not preds.shape[0] == target.shape[0]:
        raise ValueError("The `preds` and `target` should have the same first dimension.")
    if multiclass is False and ext.tmax > 1:
        raise ValueError("If you set `multiclass=False`, then `target` should not exceed 1.")
    if multiclass is False and not preds_float and ext.pmax > 1:
        raise ValueError("If you set `multiclass=False` and `preds` are integers, then `preds` should not exceed 1.")


def _check_shape_and_type_consistency(preds: Tensor, target: Tensor,
                                      ext: Optional[_Extremes] = None) -> Tuple[DataType, int]:
    """Input case and the number of classes the shapes imply (checks.py:79-137)."""
    preds_float = preds.is_floating_point()
    if preds.ndim == target.ndim:
        if preds.shape != target.shape:
            raise ValueError(
                "The `preds` and `target` should have the same shape,",
                f" got `preds` with shape={preds.shape} and `target` with shape={target.shape}.",
            )
        if preds_float and target.numel() > 0 and (ext or _Extremes(preds, target)).tmax > 1:
            raise ValueError(
                "If `preds` and `target` are of shape (N, ...) and `preds` are floats, `target` should be binary."
            )
        flat = preds.ndim == 1
        case = {(True, True): DataType.BINARY, (True, False): DataType.MULTICLASS,
                (False, True): DataType.MULTILABEL, (False, False): DataType.MULTIDIM_MULTICLASS}[(flat, preds_float)]
        return case, (preds[0].numel() if preds.numel() > 0 else 0)
    if preds.ndim == target.ndim + 1:
        if not preds_float:
            raise ValueError("If `preds` have one dimension more than `target`, `preds` should be a float tensor.")
        if preds.shape[2:] != target.shape[1:]:
            raise ValueError(
                "If `preds` have one dimension more than `target`, the shape of `preds` should be"
                " (N, C, ...), and the shape of `target` should be (N, ...)."
            )
        case = DataType.MULTICLASS if preds.ndim == 2 else DataType.MULTIDIM_MULTICLASS
        return case, (preds.shape[1] if preds.numel() > 0 else 0)
    raise ValueError(
        "Either `preds` and `target` both should have the (same) shape (N, ...), or `target` should be (N, ...)"
        " and `preds` should be (N, C, ...)."
    )


def _check_num_classes_binary(num_classes: int, multiclass: Optional[bool]) -> None:
    """checks.py:140-159."""
    if num_classes > 2:
        raise ValueError("Your data is binary, but `num_classes` is larger than 2.")
    if num_classes == 2 and not multiclass:
        raise ValueError(
            "Your data is binary and `num_classes=2`, but `multiclass` is not True."
            " Set it to True if you want to transform binary data to multi-class format."
        )
    if num_classes == 1 and multiclass:
        raise ValueError(
            "You have binary data and have set `multiclass=True`, but `num_classes` is 1."
            " Either set `multiclass=None`(default) or set `num_classes=2`"
            " to transform binary data to multi-class format."
        )


def _check_num_classes_mc(preds: Tensor, target: Tensor, num_classes: int, multiclass: Optional[bool],
                          implied_classes: int, ext: Optional[_Extremes] = None) -> None:
    """checks.py:162-192."""
    if num_classes == 1 and multiclass is not False:
        raise ValueError(
            "You have set `num_classes=1`, but predictions are integers."
            " If you want to convert (multi-dimensional) multi-class data with 2 classes"
            " to binary/multi-label, set `multiclass=False`."
        )
    if num_classes > 1:
        if multiclass is False and implied_classes != num_classes:
            raise ValueError(
                "You have set `multiclass=False`, but the implied number of classes "
                " (from shape of inputs) does not match `num_classes`. If you are trying to"
                " transform multi-dim multi-class data with 2 classes to multi-label, `num_classes`"
                " should be either None or the product of the size of extra dimensions (...)."
                " See Input Types in Metrics documentation."
            )
        if target.numel() > 0 and num_classes <= (ext or _Extremes(preds, target)).tmax:
            raise ValueError("The highest label in `target` should be smaller than `num_classes`.")
        if preds.shape != target.shape and num_classes != implied_classes:
            raise ValueError("The size of C dimension of `preds` does not match `num_classes`.")


def _check_num_classes_ml(num_classes: int, multiclass: Optional[bool], implied_classes: int) -> None:
    """checks.py:195-206."""
    if multiclass and num_classes != 2:
        raise ValueError(
            "Your have set `multiclass=True`, but `num_classes` is not equal to 2."
            " If you are trying to transform multi-label data to 2 class multi-dimensional"
            " multi-class, you should set `num_classes` to either 2 or None."
        )
    if not multiclass and num_classes != implied_classes:
        raise ValueError("The implied number of classes (from shape of inputs) does not match num_classes.")


def _check_top_k(top_k: int, case: str, implied_classes: int, multiclass: Optional[bool], preds_float: bool) -> None:
    """checks.py:209-226."""
    if case == DataType.BINARY:
        raise ValueError("You can not use `top_k` parameter with binary data.")
    if not isinstance(top_k, int) or top_k <= 0:
        raise ValueError("The `top_k` has to be an integer larger than 0.")
    if not preds_float:
        raise ValueError("You have set `top_k`, but you do not have probability predictions.")
    if multiclass is False:
        raise ValueError("If you set `multiclass=False`, you can not set `top_k`.")
    if case == DataType.MULTILABEL and multiclass:
        raise ValueError(
            "If you want to transform multi-label data to 2 class multi-dimensional"
            "multi-class data using `multiclass=True`, you can not use `top_k`."
        )
    if top_k >= implied_classes:
        raise ValueError("The `top_k` has to be strictly smaller than the `C` dimension of `preds`.")


def _check_classification_inputs(preds: Tensor, target: Tensor, threshold: float, num_classes: Optional[int],
                                 multiclass: Optional[bool], top_k: Optional[int],
                                 ignore_index: Optional[int] = None, ext: Optional[_Extremes] = None) -> DataType:
    """All of the above in the reference's order (checks.py:229-312)."""
    ext = ext or _Extremes(preds, target)
    _basic_input_validation(preds, target, threshold, multiclass, ignore_index, ext)
    case, implied_classes = _check_shape_and_type_consistency(preds, target, ext)
    if preds.shape != target.shape:
        if multiclass is False and implied_classes != 2:
            raise ValueError(
                "You have set `multiclass=False`, but have more than 2 classes in your data,"
                " based on the C dimension of `preds`."
            )
        if ext.tmax >= implied_classes:
            raise ValueError(
                "The highest label in `target` should be smaller than the size of the `C` dimension of `preds`."
            )
    if num_classes:
        if case == DataType.BINARY:
            _check_num_classes_binary(num_classes, multiclass)
        elif case in (DataType.MULTICLASS, DataType.MULTIDIM_MULTICLASS):
            _check_num_classes_mc(preds, target, num_classes, multiclass, implied_classes, ext)
        else:
            _check_num_classes_ml(num_classes, multiclass, implied_classes)
    if top_k is not None:
        _check_top_k(top_k, case, implied_classes, multiclass, preds.is_floating_point())
    return case


def _input_squeeze(preds: Tensor, target: Tensor) -> Tuple[Tensor, Tensor]:
    """Drop size-1 dimensions except the first (checks.py:301-312)."""
    if preds.shape[0] == 1:
        return preds.squeeze().unsqueeze(0), target.squeeze().unsqueeze(0)
    return preds.squeeze(), target.squeeze()


def _input_format_classification(preds: Tensor, target: Tensor, threshold: float = 0.5, top_k: Optional[int] = None,
                                 num_classes: Optional[int] = None, multiclass: Optional[bool] = None,
                                 ignore_index: Optional[int] = None) -> Tuple[Tensor, Tensor, DataType]:
    """(N, C) or (N, C, X) int32 0/1 tensors + the input case (checks.py:315-456).  Binary / multilabel scores are
    binarised with `preds >= threshold` (non-strict, unlike the task-specific `_format` functions)."""
    preds, target = _input_squeeze(preds, target)
    if preds.dtype == torch.float16:
        preds = preds.float()
    ext = _Extremes(preds, target)
    case = _check_classification_inputs(preds, target, threshold=threshold, num_classes=num_classes,
                                        multiclass=multiclass, top_k=top_k, ignore_index=ignore_index, ext=ext)
    label_like = case in (DataType.MULTICLASS, DataType.MULTIDIM_MULTICLASS)
    if not label_like and not top_k:
        preds = (preds >= threshold).int()
        num_classes = num_classes if not multiclass else 2
    if case == DataType.MULTILABEL and top_k:
        preds = select_topk(preds, top_k)
    if label_like or multiclass:
        if preds.is_floating_point():
            num_classes = preds.shape[1]
            preds = select_topk(preds, top_k or 1)
        else:
            if not num_classes:
                # after binarisation preds are 0/1: their own extremes, not the raw scores', decide here
                pmax = ext.pmax if label_like else float(preds.max()) if preds.numel() else 0.0
                num_classes = int(max(pmax, ext.tmax) + 1)
            preds = to_onehot(preds, max(2, num_classes))
        target = to_onehot(target, max(2, num_classes))
        if multiclass is False:
            preds, target = preds[:, 1, ...], target[:, 1, ...]
    if not _check_for_empty_tensors(preds, target):
        if (label_like and multiclass is not False) or multiclass:
            target = target.reshape(target.shape[0], target.shape[1], -1)
            preds = preds.reshape(preds.shape[0], preds.shape[1], -1)
        else:
            target = target.reshape(target.shape[0], -1)
            preds = preds.reshape(preds.shape[0], -1)
    if preds.ndim > 2:
        preds, target = preds.squeeze(-1), target.squeeze(-1)
    return preds.int(), target.int(), case
