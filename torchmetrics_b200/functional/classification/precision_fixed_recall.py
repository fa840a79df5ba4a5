"""precision at fixed recall (reference: functional/classification/precision_fixed_recall.py); see `_operating_point.py`."""
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

_F = _op.PRECISION_AT_RECALL
_Thr = Optional[Union[int, List[float], Tensor]]

_binary_precision_at_fixed_recall_arg_validation = _F.binary_arg_validation
_multiclass_precision_at_fixed_recall_arg_validation = _F.multiclass_arg_validation
_multilabel_precision_at_fixed_recall_arg_validation = _F.multilabel_arg_validation


def binary_precision_at_fixed_recall(preds: Tensor, target: Tensor, min_recall: float, thresholds: _Thr = None,
           ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.binary(preds, target, min_recall, thresholds, ignore_index, validate_args)


def multiclass_precision_at_fixed_recall(preds: Tensor, target: Tensor, num_classes: int, min_recall: float, thresholds: _Thr = None,
               ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.multiclass(preds, target, num_classes, min_recall, thresholds, ignore_index, validate_args)


def multilabel_precision_at_fixed_recall(preds: Tensor, target: Tensor, num_labels: int, min_recall: float, thresholds: _Thr = None,
               ignore_index: Optional[int] = None, validate_args: bool = True) -> Tuple[Tensor, Tensor]:
    return _F.multilabel(preds, target, num_labels, min_recall, thresholds, ignore_index, validate_args)


def precision_at_fixed_recall(preds: Tensor, target: Tensor, task: Literal["binary", "multiclass", "multilabel"], min_recall: float,
    thresholds: _Thr = None, num_classes: Optional[int] = None, num_labels: Optional[int] = None,
    ignore_index: Optional[int] = None, validate_args: bool = True):
    return _F.dispatch(preds, target, task, min_recall, thresholds, num_classes, num_labels, ignore_index,
                       validate_args)
