"""sensitivity at specificity (reference: functional/classification/sensitivity_specificity.py); see `_operating_point.py`."""
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

_F = _op.SENSITIVITY_AT_SPECIFICITY
_Thr = Optional[Union[int, List[float], Tensor]]

_binary_sensitivity_at_specificity_arg_validation = _F.binary_arg_validation
_multiclass_sensitivity_at_specificity_arg_validation = _F.multiclass_arg_validation
_multilabel_sensitivity_at_specificity_arg_validation = _F.multilabel_arg_validation


def _binary_sensitivity_at_specificity_compute(state, thresholds, min_specificity: float, pos_label: int = 1):
    return _F.binary_compute(state, thresholds, min_specificity, pos_label)


def _multiclass_sensitivity_at_specificity_compute(state, num_classes: int, thresholds, min_specificity: float):
    return _F.multiclass_compute(state, num_classes, thresholds, min_specificity)


def _multilabel_sensitivity_at_specificity_compute(state, num_labels: int, thresholds, ignore_index: Optional[int],
                                                   min_specificity: float):
    return _F.multilabel_compute(state, num_labels, thresholds, ignore_index, min_specificity)


def binary_sensitivity_at_specificity(preds: Tensor, target: Tensor, min_specificity: float, thresholds: _Thr = None,
           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.binary(preds, target, min_specificity, thresholds, ignore_index, validate_args)


def multiclass_sensitivity_at_specificity(preds: Tensor, target: Tensor, num_classes: int, min_specificity: float, thresholds: _Thr = None,
               ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.multiclass(preds, target, num_classes, min_specificity, thresholds, ignore_index, validate_args)


def multilabel_sensitivity_at_specificity(preds: Tensor, target: Tensor, num_labels: int, min_specificity: float, thresholds: _Thr = None,
               ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.multilabel(preds, target, num_labels, min_specificity, thresholds, ignore_index, validate_args)


def sensitivity_at_specificity(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"], min_specificity: float,
    thresholds: _Thr = None, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
    ignore_index: Optional[int] = None, validate_args: bool = True):
    return _F.dispatch(preds, target, task, min_specificity, thresholds, num_classes, num_labels, ignore_index,
                       validate_args)
