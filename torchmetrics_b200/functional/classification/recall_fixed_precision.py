"""recall at fixed precision (reference: functional/classification/recall_fixed_precision.py); see `_operating_point.py`."""
from typing import Callable, List, Optional, Tuple, Union

from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.functional.classification import _operating_point as _op
from torchmetrics_b200.functional.classification._operating_point import (  # noqa: F401
    _convert_fpr_to_specificity,
    _lexargmax,
    _precision_at_recall,
    _recall_at_precision,
    _sensitivity_at_specificity,
    _specificity_at_sensitivity,
)

_F = _op.RECALL_AT_PRECISION
_Thr = Optional[Union[int, List[float], Tensor]]

_binary_recall_at_fixed_precision_arg_validation = _F.binary_arg_validation
_multiclass_recall_at_fixed_precision_arg_validation = _F.multiclass_arg_validation
_multilabel_recall_at_fixed_precision_arg_validation = _F.multilabel_arg_validation


def _binary_recall_at_fixed_precision_compute(state, thresholds, min_precision: float, pos_label: int = 1,
                                              reduce_fn: Callable = _recall_at_precision):
    return _F.binary_compute(state, thresholds, min_precision, pos_label, reduce_fn)


def _multiclass_recall_at_fixed_precision_arg_compute(state, num_classes: int, thresholds, min_precision: float,
                                                      reduce_fn: Callable = _recall_at_precision):
    return _F.multiclass_compute(state, num_classes, thresholds, min_precision, reduce_fn)


def _multilabel_recall_at_fixed_precision_arg_compute(state, num_labels: int, thresholds, ignore_index: Optional[int],
                                                      min_precision: float,
                                                      reduce_fn: Callable = _recall_at_precision):
    return _F.multilabel_compute(state, num_labels, thresholds, ignore_index, min_precision, reduce_fn)


def binary_recall_at_fixed_precision(preds: Tensor, target: Tensor, min_precision: float, thresholds: _Thr = None,
           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.binary(preds, target, min_precision, thresholds, ignore_index, validate_args)


def multiclass_recall_at_fixed_precision(preds: Tensor, target: Tensor, num_classes: int, min_precision: float, thresholds: _Thr = None,
               ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.multiclass(preds, target, num_classes, min_precision, thresholds, ignore_index, validate_args)


def multilabel_recall_at_fixed_precision(preds: Tensor, target: Tensor, num_labels: int, min_precision: float, thresholds: _Thr = None,
               ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.multilabel(preds, target, num_labels, min_precision, thresholds, ignore_index, validate_args)


def recall_at_fixed_precision(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"], min_precision: float,
    thresholds: _Thr = None, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
    ignore_index: Optional[int] = None, validate_args: bool = True):
    return _F.dispatch(preds, target, task, min_precision, thresholds, num_classes, num_labels, ignore_index,
                       validate_args)
