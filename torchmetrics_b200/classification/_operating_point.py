"""Stateful operating-point metrics (reference: classification/{recall_fixed_precision,precision_fixed_recall,
sensitivity_specificity,specificity_sensitivity}.py).  State and `update` are the precision-recall-curve classes'
(one K3 pass per batch, or cat states for the exact curve); only `compute` differs.  `make_classes` builds the
Binary / Multiclass / Multilabel classes and the task wrapper of one family."""
from typing import Any, Optional, Tuple

from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_b200.classification.precision_recall_curve import (
    BinaryPrecisionRecallCurve,
    MulticlassPrecisionRecallCurve,
    MultilabelPrecisionRecallCurve,
    _Thresholds,
)
from torchmetrics_b200.functional.classification._operating_point import Family
from torchmetrics_b200.metric import Metric
from torchmetrics_b200.utilities.enums import ClassificationTask


def make_classes(family: Family, stem: str, attr: str, module: str):
    """`stem` e.g. "RecallAtFixedPrecision"; `attr` the constructor argument / attribute holding the minimum."""
    common = dict(is_differentiable=False, higher_is_better=None, full_state_update=False, plot_lower_bound=0.0,
                  plot_upper_bound=1.0, __module__=module)

    def binary_init(self, *args: Any, thresholds: _Thresholds = None, ignore_index: Optional[int] = None,
                    validate_args: bool = True, **kwargs: Any) -> None:
        minimum = args[0] if args else kwargs.pop(attr)
        BinaryPrecisionRecallCurve.__init__(self, thresholds, ignore_index, validate_args=False, **kwargs)
        if validate_args:
            family.binary_arg_validation(minimum, thresholds, ignore_index)
        self.validate_args = validate_args
        setattr(self, attr, minimum)

    def binary_compute(self) -> Tuple[Tensor, Tensor]:
        return family.binary_compute(self._state(), self.thresholds, getattr(self, attr))

    def multiclass_init(self, num_classes: int, *args: Any, thresholds: _Thresholds = None,
                        ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> None:
        minimum = args[0] if args else kwargs.pop(attr)
        MulticlassPrecisionRecallCurve.__init__(self, num_classes=num_classes, thresholds=thresholds,
                                                ignore_index=ignore_index, validate_args=False, **kwargs)
        if validate_args:
            family.multiclass_arg_validation(num_classes, minimum, thresholds, ignore_index)
        self.validate_args = validate_args
        setattr(self, attr, minimum)

    def multiclass_compute(self) -> Tuple[Tensor, Tensor]:
        return family.multiclass_compute(self._state(), self.num_classes, self.thresholds, getattr(self, attr))

    def multilabel_init(self, num_labels: int, *args: Any, thresholds: _Thresholds = None,
                        ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> None:
        minimum = args[0] if args else kwargs.pop(attr)
        MultilabelPrecisionRecallCurve.__init__(self, num_labels=num_labels, thresholds=thresholds,
                                                ignore_index=ignore_index, validate_args=False, **kwargs)
        if validate_args:
            family.multilabel_arg_validation(num_labels, minimum, thresholds, ignore_index)
        self.validate_args = validate_args
        setattr(self, attr, minimum)

    def multilabel_compute(self) -> Tuple[Tensor, Tensor]:
        return family.multilabel_compute(self._state(), self.num_labels, self.thresholds, self.ignore_index,
                                         getattr(self, attr))

    binary = type("Binary" + stem, (BinaryPrecisionRecallCurve,),
                  dict(common, __init__=binary_init, compute=binary_compute))
    multiclass = type("Multiclass" + stem, (MulticlassPrecisionRecallCurve,),
                      dict(common, __init__=multiclass_init, compute=multiclass_compute, plot_legend_name="Class"))
    multilabel = type("Multilabel" + stem, (MultilabelPrecisionRecallCurve,),
                      dict(common, __init__=multilabel_init, compute=multilabel_compute, plot_legend_name="Label"))

    def wrapper_new(cls, task: Literal["binary", "multiclass", "multilabel"], *args: Any,
                    thresholds: _Thresholds = None, num_classes: Optional[int] = None,
                    num_labels: Optional[int] = None, ignore_index: Optional[int] = None, validate_args: bool = True,
                    **kwargs: Any) -> Metric:
        minimum = args[0] if args else kwargs.pop(attr)
        task = ClassificationTask.from_str(task)
        kw = dict(kwargs, thresholds=thresholds, ignore_index=ignore_index, validate_args=validate_args)
        if task == ClassificationTask.BINARY:
            return binary(minimum, **kw)
        if task == ClassificationTask.MULTICLASS:
            if not isinstance(num_classes, int):
                raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
            return multiclass(num_classes, minimum, **kw)
        if task == ClassificationTask.MULTILABEL:
            if not isinstance(num_labels, int):
                raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
            return multilabel(num_labels, minimum, **kw)
        raise ValueError(f"Task {task} not supported!")

    wrapper = type(stem, (_ClassificationTaskWrapper,), {"__new__": wrapper_new, "__module__": module})
    return binary, multiclass, multilabel, wrapper
