"""specificity at sensitivity (reference: functional/classification/specificity_sensitivity.py); see `_operating_point.py`."""
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

_F = _op.SPECIFICITY_AT_SENSITIVITY
_Thr = Optional[Union[int, List[float], Tensor]]

_binary_specificity_at_sensitivity_arg_validation = _F.binary_arg_validation
_multiclass_specificity_at_sensitivity_arg_validation = _F.multiclass_arg_validation
_multilabel_specificity_at_sensitivity_arg_validation = _F.multilabel_arg_validation


def _binary_specificity_at_sensitivity_compute(state, thresholds, min_sensitivity: float, pos_label: int = 1):
    return _F.binary_compute(state, thresholds, min_sensitivity, pos_label)


def _multiclass_specificity_at_sensitivity_compute(state, num_classes: int, thresholds, min_sensitivity: float):
    return _F.multiclass_compute(state, num_classes, thresholds, min_sensitivity)


def _multilabel_specificity_at_sensitivity_compute(state, num_labels: int, thresholds, ignore_index: Optional[int],
                                                   min_sensitivity: float):
    return _F.multilabel_compute(state, num_labels, thresholds, ignore_index, min_sensitivity)


def binary_specificity_at_sensitivity(preds: Tensor, target: Tensor, min_sensitivity: float, thresholds: _Thr = None,
           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.binary(preds, target, min_sensitivity, thresholds, ignore_index, validate_args)


def multiclass_specificity_at_sensitivity(preds: Tensor, target: Tensor, num_classes: int, min_sensitivity: float, thresholds: _Thr = None,
               ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.multiclass(preds, target, num_classes, min_sensitivity, thresholds, ignore_index, validate_args)


def multilabel_specificity_at_sensitivity(preds: Tensor, target: Tensor, num_labels: int, min_sensitivity: float, thresholds: _Thr = None,
               ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.multilabel(preds, target, num_labels, min_sensitivity, thresholds, ignore_index, validate_args)


def specificity_at_sensitivity(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"], min_sensitivity: float,
    thresholds: _Thr = None, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
    ignore_index: Optional[int] = None, validate_args: bool = True):
    return _F.dispatch(preds, target, task, min_sensitivity, thresholds, num_classes, num_labels, ignore_index,
                       validate_args)
